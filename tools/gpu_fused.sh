#!/bin/bash
# Quick GPU pass for the fused modulate+AWGN+bank kernel: its parity tests, then fused vs four-kernel timings.
# Usage (on the GPU box, from the repo root):  bash tools/gpu_fused.sh [tag]
TAG=${1:-r1p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider -x -k "fused or raw_bank or noiseless" > $OUT/pytest_fused.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_fused.log
tail -3 $OUT/pytest_fused.log
timeout 300 python tools/kbench.py --fused --frames 65536 --reps 5 >> $OUT/kbench_fused.log 2>&1
timeout 300 python tools/kbench.py --frames 65536 --reps 5 >> $OUT/kbench_fused.log 2>&1
cat $OUT/kbench_fused.log
