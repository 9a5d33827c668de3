#!/bin/bash
# 1-GPU visit: racecheck of the P3 kernels on small cases, then C5 at full size (1 step).
O=gpurun_out; mkdir -p $O
timeout -s KILL 600 compute-sanitizer --tool racecheck --print-limit 5 python -m pytest tests/test_p3.py -m gpu -x -q -k "matches_oracle or max_color" > $O/p3_racecheck.log 2>&1; echo "racecheck exit $?"
tail -5 $O/p3_racecheck.log
timeout -s KILL 900 python bench.py --workload c5 --steps 1 --warmup 1 --no-extra > $O/c5_n1.json 2> $O/c5_n1.err; echo "c5 exit $?"; tail -2 $O/c5_n1.err
cut -c1-400 $O/c5_n1.json
