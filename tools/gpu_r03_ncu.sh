#!/bin/bash
# Round 3 ncu captures at the BENCHMARKED configuration ("traffic" of the timed launch, C5 kernel, k_primary).
#  1. C4 16384^2, first 16 Mi-pixel tile: k_shade levels 0..6 (--set full) and k_primary -> DRAM bytes per launch
#  2. launch list of one C4 frame (gpu__time_duration only): share of each kernel class in the step
#  3. C5 at 4096^2: k_shade levels 0 and 1 of the current kernel (ball pre-test included)
O=gpurun_out; mkdir -p $O
A4="--steps 1 --warmup 0 --no-cpu --no-extra"
timeout 300 python bench.py $A4 > $O/r03_ncu_plain_c4.json 2> $O/r03_ncu_plain_c4.err || { tail -5 $O/r03_ncu_plain_c4.err; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -c 7 -o $O/r03_k_shade_c4 -f python bench.py $A4 > $O/r03_ncu_c4_shade.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_primary -c 1 -o $O/r03_k_primary_c4 -f python bench.py $A4 > $O/r03_ncu_c4_primary.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r03_ncu_launches_c4.csv python bench.py $A4 > $O/r03_ncu_c4_list.log 2>&1
A5="--workload c5 --size 4096 --steps 1 --warmup 1 --no-cpu --no-extra"
timeout 300 python bench.py $A5 > $O/r03_ncu_plain_c5.json 2> $O/r03_ncu_plain_c5.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 2 -o $O/r03_k_shade_c5 -f python bench.py $A5 > $O/r03_ncu_c5_shade.log 2>&1
ls -la $O | grep r03_
