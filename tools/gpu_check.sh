#!/bin/bash
# One GPU-box visit: GPU parity tests, the default bench, an ncu launch list and one full capture of k_shade.
# Usage (from the repo root, under gpurun): bash tools/gpu_check.sh [tag]
TAG=${1:-run}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit $?"
timeout 300 python bench.py --size 2048 --steps 1 --warmup 1 --no-cpu --no-extra > $O/${TAG}_bench2048.json 2>> $O/${TAG}_bench.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --size 2048 --steps 1 --warmup 1 --no-cpu --no-extra > $O/${TAG}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 2 -o $O/${TAG}_k_shade -f \
  python bench.py --size 2048 --steps 1 --warmup 1 --no-cpu --no-extra > $O/${TAG}_ncu2.log 2>&1
ls -la $O
