#!/bin/bash
# Round 2, visit A: scan-shape microbenchmark, one-shot CLI stage trace (C1-C3) next to the reference CLI,
# GPU tests of the refactored launcher, quick bench.
O=gpurun_out; mkdir -p $O; S=tests/golden/scenes
M=raytracing-programs_b200/bin/raytrace; R=oracle/_ref/raytrace_ref
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/r02a_smi.txt
grep -m1 "model name" /proc/cpuinfo >> $O/r02a_smi.txt; nproc >> $O/r02a_smi.txt
timeout 300 ./raytracing-programs_b200/bin/scan_shapes > $O/r02_scan_shapes.json 2>&1
: > $O/r02a_cli.txt
for cfg in "1000 1000 ref_simple" "3840 2160 ref_reflection" "7680 4320 ref_quadric"; do
  set -- $cfg
  cp $S/$3.scene /tmp/s.scene; chmod u+w /tmp/s.scene
  for k in 1 2 3; do
    echo "== ours $3 $1x$2 run $k" >> $O/r02a_cli.txt
    /usr/bin/time -f "wall %e s" env RT_TRACE=1 RT_VERBOSE=1 timeout 600 $M $1 $2 /tmp/s.scene /tmp/our_$3.ppm >> $O/r02a_cli.txt 2>&1
  done
  echo "== reference CLI $3 $1x$2" >> $O/r02a_cli.txt
  /usr/bin/time -f "wall %e s" timeout 900 $R $1 $2 /tmp/s.scene /tmp/ref_$3.ppm >> $O/r02a_cli.txt 2>&1
  cmp /tmp/our_$3.ppm /tmp/ref_$3.ppm >> $O/r02a_cli.txt 2>&1 && echo "files identical" >> $O/r02a_cli.txt
done
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02a_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02a_pytest.log
tail -3 $O/r02a_pytest.log
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > $O/r02a_c4.json 2> $O/r02a_c4.err; echo "bench exit $?"
tail -c 300 $O/r02a_c4.json
