#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_shoot_mode.py tests/test_gpu_parity.py tests/test_depth.py -m gpu -x -q > $O/r02i_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02i_pytest.log
tail -3 $O/r02i_pytest.log
bash tools/gpu_ncu_kshade.sh r02i_c5 --workload c5
