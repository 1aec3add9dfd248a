#!/bin/bash
TAG=${1:-r2k}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log; tail -3 $OUT/pytest_gpu.log
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform soqpsk-b >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
