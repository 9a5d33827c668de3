#!/bin/bash
# 1-GPU visit: GPU parity suite + smoke().
timeout -s KILL 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
