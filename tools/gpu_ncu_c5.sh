#!/bin/bash
O=gpurun_out; mkdir -p $O
python bench.py --workload c5 --size 1536 --steps 1 --warmup 1 --no-cpu --no-extra > $O/c5ncu_plain.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 1 -o $O/c5_k_shade -f \
  python bench.py --workload c5 --size 1536 --steps 1 --warmup 1 --no-cpu --no-extra > $O/c5ncu.log 2>&1
tail -2 $O/c5ncu.log | cut -c1-200
