#!/bin/bash
# Round 2, visit B: what caps the reject-scan loop (ncu pipe + stall metrics of the microbenchmark shapes),
# one-shot CLI stage trace (C1-C3) next to the reference CLI.
O=gpurun_out; mkdir -p $O; S=tests/golden/scenes
M=raytracing-programs_b200/bin/raytrace; R=oracle/_ref/raytrace_ref; B=./raytracing-programs_b200/bin/scan_shapes
SCAN_SHAPES_SHORT=1 timeout 120 $B > $O/r02b_scan_variants.json 2>&1
cat $O/r02b_scan_variants.json
MET=smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__cycles_elapsed.max,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio
for w in A0 A1 A2 A3 B C; do
  timeout 300 ncu --metrics $MET --clock-control none -k regex:k_ -s 1 -c 1 --csv --log-file $O/r02b_ncu_$w.csv $B $w 4 > /dev/null 2>&1
done
: > $O/r02b_cli.txt
for cfg in "1000 1000 ref_simple" "3840 2160 ref_reflection" "7680 4320 ref_quadric"; do
  set -- $cfg
  cp $S/$3.scene /tmp/s.scene; chmod u+w /tmp/s.scene
  for k in 1 2 3; do
    echo "== ours $3 $1x$2 run $k" >> $O/r02b_cli.txt
    { time RT_TRACE=1 RT_VERBOSE=1 timeout 600 $M $1 $2 /tmp/s.scene /tmp/our_$3.ppm ; } >> $O/r02b_cli.txt 2>&1
  done
  echo "== reference CLI $3 $1x$2" >> $O/r02b_cli.txt
  { time timeout 900 $R $1 $2 /tmp/s.scene /tmp/ref_$3.ppm ; } >> $O/r02b_cli.txt 2>&1
  cmp /tmp/our_$3.ppm /tmp/ref_$3.ppm >> $O/r02b_cli.txt 2>&1 && echo "files identical" >> $O/r02b_cli.txt
done
grep -E "^==|Time|B200|real|identical|rt_trace" $O/r02b_cli.txt | head -150
