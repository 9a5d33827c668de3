#!/bin/bash
# 8-GPU visit: C4 at N = 8, 4, 2 and C5 at N = 8 (the driver's own scaling run uses the same launch line).
O=gpurun_out; mkdir -p $O
run() { # N workload steps warmup tag
  timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + $1)) \
    bench.py --gpus $1 --workload $2 --steps $3 --warmup $4 --no-cpu > $O/$5.json 2> $O/$5.err; echo "$5 exit $?"; cut -c1-200 $O/$5.json; tail -2 $O/$5.err
}
nvidia-smi -L | head -8
run 8 c4 3 3 scale_c4_n8
run 8 c5 2 1 scale_c5_n8
run 4 c4 3 3 scale_c4_n4
run 2 c4 3 3 scale_c4_n2
RT_ACCEL=1 run 8 c4 3 3 scale_c4_n8_accel
