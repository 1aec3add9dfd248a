#!/bin/bash
TAG=${1:-r2m}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -s -k "ber_waterfall" > $OUT/pytest_ber.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_ber.log; tail -50 $OUT/pytest_ber.log
