#!/bin/bash
TAG=$1; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
bash scripts/gpu_opt.sh $TAG "usum_period=0 usum_period=64" "c2 c4 c3"
