#!/bin/bash
# Context comparator: the reference's own CUDA variant (oracle/_ref/raytrace_refcuda, unmodified sources,
# sm_100a) against this repo's CLI on C1-C3.  Kernel time of the reference from ncu (its own timer
# includes the fprintf P3 write); ours from RT_VERBOSE.  Also a pixel comparison of the two images.
O=gpurun_out; mkdir -p $O; S=tests/golden/scenes
R=oracle/_ref/raytrace_refcuda; M=raytracing-programs_b200/bin/raytrace
: > $O/refcuda.txt
for cfg in "1000 1000 ref_simple" "3840 2160 ref_reflection" "7680 4320 ref_quadric"; do
  set -- $cfg
  echo "== $3 $1x$2" >> $O/refcuda.txt
  cp $S/$3.scene /tmp/s.scene; chmod u+w /tmp/s.scene
  timeout 600 $R $1 $2 /tmp/s.scene /tmp/ref_$3.ppm >> $O/refcuda.txt 2>&1
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv $R $1 $2 /tmp/s.scene /dev/null 2>&1 | grep -E "raytrace_engine|warm" | awk -F'","' '{print $5, $(NF)}' >> $O/refcuda.txt
  RT_VERBOSE=1 timeout 600 $M $1 $2 /tmp/s.scene /tmp/our_$3.ppm >> $O/refcuda.txt 2>&1
  RT_VERBOSE=1 timeout 600 $M $1 $2 /tmp/s.scene /tmp/our_$3.ppm >> $O/refcuda.txt 2>&1
  python - "$3" >> $O/refcuda.txt 2>&1 <<'PY'
import sys, numpy as np
def rd(p):
    d = open(p, 'rb').read().split(b'\n')
    return np.array(d[3:-1], dtype=np.int32)
a, b = rd(f'/tmp/ref_{sys.argv[1]}.ppm'), rd(f'/tmp/our_{sys.argv[1]}.ppm')
n = min(a.size, b.size); a = a[:n].reshape(-1, 3); b = b[:n].reshape(-1, 3)
d = np.abs(a - b).max(axis=1)
print(f"pixels {d.size} identical {(d == 0).mean() * 100:.4f} % max diff {d.max()} (reference CUDA variant vs this repo)")
PY
done
cat $O/refcuda.txt
