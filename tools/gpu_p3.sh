#!/bin/bash
# GPU visit for the P3 writer: tests, bench leg, ncu of k_p3_format.
TAG=${1:-p3}
O=gpurun_out; mkdir -p $O
timeout -s KILL 600 python -m pytest tests/test_p3.py -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -15 $O/${TAG}_pytest.log
grep -q "pytest exit 0" $O/${TAG}_pytest.log || exit 1
timeout -s KILL 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"; tail -3 $O/${TAG}_bench.err
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:k_p3_ -s 6 -c 3 -o $O/${TAG}_k_p3 -f \
  python bench.py --size 8192 --steps 1 --warmup 1 --no-cpu > $O/${TAG}_ncu.log 2>&1
tail -2 $O/${TAG}_ncu.log | cut -c1-300
