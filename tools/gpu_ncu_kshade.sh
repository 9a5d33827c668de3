#!/bin/bash
# ncu of k_shade (levels 0 and 1 of a 2048^2 C4 frame): full set + launch list.  usage: gpu_ncu_kshade.sh tag [workload args]
TAG=${1:-ncu}; shift
O=gpurun_out; mkdir -p $O
ARGS="--size 2048 --steps 1 --warmup 1 --no-cpu --no-extra $@"
timeout 300 python bench.py $ARGS > $O/${TAG}_bench2048.json 2> $O/${TAG}_bench.err || { tail -5 $O/${TAG}_bench.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py $ARGS > $O/${TAG}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 2 -o $O/${TAG}_k_shade -f python bench.py $ARGS > $O/${TAG}_ncu2.log 2>&1
ls -la $O | grep $TAG
