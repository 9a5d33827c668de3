#!/bin/bash
# Last GPU visit of round 2: the default bench and the reference arm as the driver runs them, C5 N=1 with the final build,
# ncu launch list + capture of the final k_shade, then the GPU tests.
O=gpurun_out; mkdir -p $O
timeout 200 python bench.py > $O/r02_final_bench.json 2> $O/r02_final_bench.err; echo "bench exit $?"
timeout 60 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_final_ref.json 2> $O/r02_final_ref.err; echo "ref exit $?"
timeout 150 python bench.py --workload c5 --steps 1 --warmup 1 --no-cpu --no-extra > $O/r02_final_c5_n1.json 2> $O/r02_final_c5_n1.err; echo "c5 exit $?"
ARGS="--size 2048 --steps 1 --warmup 1 --no-cpu --no-extra"
timeout 60 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_final_launches.csv python bench.py $ARGS > $O/r02_final_ncu1.log 2>&1
timeout 90 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 2 -o $O/r02_final_k_shade -f python bench.py $ARGS > $O/r02_final_ncu2.log 2>&1
timeout 170 python -m pytest tests -m gpu -q -x > $O/r02_final_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02_final_pytest.log
tail -3 $O/r02_final_pytest.log
python - <<PY
import json
for f in ("r02_final_bench", "r02_final_c5_n1"):
    try:
        b = json.loads([l for l in open(f"$O/{f}.json") if l.startswith("{")][-1])
        print(f, round(b["value"], 1), "Mrays/s", round(b["ms_per_step"], 2), "ms e2e", round(b["e2e"]["value"], 1), "frac", round(b["roofline"]["frac"], 4), (b.get("parity") or {}).get("identical_frac"))
    except Exception as e:
        print(f, "ERR", e)
PY
