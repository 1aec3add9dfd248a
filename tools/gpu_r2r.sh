#!/bin/bash
TAG=${1:-r2r}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "cpm or generic" > $OUT/pytest_cpm.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_cpm.log
for i in 1 2; do timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform pcmfm >> $OUT/kbench.log 2>&1; done
timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform pcmfm --extra '{"tune": 1}' >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
