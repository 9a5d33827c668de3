#!/bin/bash
# Round-2 scaling visits.  usage: gpu_scale2.sh <1|2|8>   (the number of GPUs of the box)
O=gpurun_out; mkdir -p $O
run() { # N workload steps warmup tag [extra bench args]
  local N=$1 W=$2 S=$3 U=$4 T=$5; shift 5
  if [ $N -eq 1 ]; then
    timeout -s KILL 1500 python bench.py --gpus 1 --workload $W --steps $S --warmup $U "$@" > $O/$T.json 2> $O/$T.err
  else
    timeout -s KILL 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --workload $W --steps $S --warmup $U "$@" > $O/$T.json 2> $O/$T.err
  fi
  echo "$T exit $?"; python - <<PY
import json
try:
    b = json.loads([l for l in open("$O/$T.json") if l.startswith("{")][-1])
    print("  ", "$T", "N", b["n_gpus"], round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 1), "ms  e2e", round(b["e2e"]["value"], 1), "frac", round(b["roofline"]["frac"], 3), "parity", (b.get("parity") or {}).get("identical_frac"), b.get("library_multi_gpu_selftest"))
except Exception as e:
    print("  ", "$T", "ERR", e, open("$O/$T.err").read()[-500:])
PY
}
nvidia-smi -L | head -8
case $1 in
1)
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02k_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02k_pytest.log; tail -3 $O/r02k_pytest.log
  run 1 c5 2 1 r02_scale_c5_n1
  RT_ACCEL=1 run 1 c5 2 1 r02_scale_c5_n1_accel --no-extra
  ;;
2)
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02k_pytest_2gpu.log 2>&1; echo "pytest exit $?" >> $O/r02k_pytest_2gpu.log; tail -3 $O/r02k_pytest_2gpu.log
  run 2 c4 3 3 r02_scale_c4_n2
  run 2 c5 2 1 r02_scale_c5_n2
  RT_ACCEL=1 run 2 c5 2 1 r02_scale_c5_n2_accel
  ;;
8)
  run 8 c5 2 1 r02_scale_c5_n8
  RT_ACCEL=1 run 8 c5 2 1 r02_scale_c5_n8_accel
  run 4 c5 2 1 r02_scale_c5_n4
  RT_ACCEL=1 run 4 c5 2 1 r02_scale_c5_n4_accel
  run 8 c4 3 3 r02_scale_c4_n8
  ;;
esac
