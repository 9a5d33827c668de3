#!/bin/bash
# Round 3, after the tile default moved to 64 Mi pixels: DRAM bytes and pipe figures of EVERY k_shade launch of one
# benchmarked C4 16384^2 frame (4 tiles x 7 levels = 28 launches; a short metric list, few replay passes), plus one
# --set full capture (source page) of the two heavy launches of tile 1 (levels 0 and 1).
O=gpurun_out; mkdir -p $O
A4="--steps 1 --warmup 0 --no-cpu --no-extra"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct
timeout 300 python bench.py $A4 > $O/r03c_plain_c4.json 2> $O/r03c_plain_c4.err || { tail -5 $O/r03c_plain_c4.err; exit 1; }
timeout 1200 ncu --metrics $M --clock-control none -k regex:k_shade -c 28 --csv --log-file $O/r03c_k_shade_frame_c4.csv python bench.py $A4 > $O/r03c_ncu_frame.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 2 -o $O/r03c_k_shade_c4_tile1 -f python bench.py $A4 > $O/r03c_ncu_full.log 2>&1
ls -la $O | grep r03c_
