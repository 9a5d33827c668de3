#!/bin/bash
# Round-end style visit: reference arm, own arm (default flags), C5 at full size.
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/final_ref.json 2> $O/final_ref.err; echo "ref exit $?"
timeout 900 python bench.py > $O/final_c4.json 2> $O/final_c4.err; echo "c4 exit $?"; tail -2 $O/final_c4.err
timeout 900 python bench.py --workload c5 --steps 1 --warmup 1 --no-extra > $O/final_c5.json 2> $O/final_c5.err; echo "c5 exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/final_smoke.log
cut -c1-300 $O/final_ref.json
