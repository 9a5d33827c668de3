#!/bin/bash
O=gpurun_out; mkdir -p $O
export RT_ACCEL=1
python bench.py --size 2048 --steps 1 --warmup 1 --no-cpu --no-extra > $O/accelncu_plain.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 1 -o $O/accel_k -f \
  python bench.py --size 2048 --steps 1 --warmup 1 --no-cpu --no-extra > $O/accelncu.log 2>&1
tail -2 $O/accelncu.log | cut -c1-200
