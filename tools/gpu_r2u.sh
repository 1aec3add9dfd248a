#!/bin/bash
TAG=${1:-r2u}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dist.py -m gpu -q -p no:cacheprovider -s -k "cpm or generic or two_gpu" > $OUT/pytest_cpm.log 2>&1; echo "pytest exit $?"; grep -v "float32 flips" $OUT/pytest_cpm.log | tail -8
for wf in pcmfm multih; do for i in 1 2; do timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform $wf >> $OUT/kbench.log 2>&1; done; done
cat $OUT/kbench.log
