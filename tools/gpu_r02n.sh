#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/r02n_pytest.log 2>&1; echo "pytest exit $?" >> $O/r02n_pytest.log
tail -3 $O/r02n_pytest.log
bash tools/gpu_r02m.sh
