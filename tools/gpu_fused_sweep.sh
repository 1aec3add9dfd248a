#!/bin/bash
# Sweep of the fused kernel's launch variants and first-wave stagger (cycles).
TAG=${1:-r1p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in ${VARIANTS:-1 3}; do
  for st in ${STAGGERS:-0 2500}; do
    echo "variant $v stagger $st" >> $OUT/sweep.log
    WF_FUSE_VARIANT=$v WF_FUSE_STAGGER=$st timeout 300 python tools/kbench.py --fused --frames 65536 --reps 5 2>&1 | grep -o '"modulate_awgn_bank": {[^}]*}' >> $OUT/sweep.log
  done
done
cat $OUT/sweep.log
