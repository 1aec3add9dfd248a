#!/bin/bash
# r2i: ARTM detector occupancy variants (tune high byte = resident blocks per SM asked of the compiler)
TAG=${1:-r2i}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "mil or cpm" > $OUT/pytest_sel.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_sel.log
for t in 4 1540 1796 2052; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform multih --extra "{\"tune\": $t}" >> $OUT/kbench.log 2>&1
done
cat $OUT/kbench.log
