#!/bin/bash
# Quick GPU pass for the fused modulate+bank kernel: its parity tests, then fused vs four-kernel timings.
TAG=${1:-r1p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in ${VARIANTS:-2}; do
  WF_FUSE_VARIANT=$v timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider -x -k "fused or raw_bank or noiseless" > $OUT/pytest_fused_v$v.log 2>&1
  echo "variant $v pytest exit $?" | tee -a $OUT/pytest_fused_v$v.log
  tail -3 $OUT/pytest_fused_v$v.log
  echo "variant $v" >> $OUT/kbench_fused.log
  WF_FUSE_VARIANT=$v timeout 300 python tools/kbench.py --fused --frames 65536 --reps 5 >> $OUT/kbench_fused.log 2>&1
done
timeout 300 python tools/kbench.py --frames 65536 --reps 5  >> $OUT/kbench_fused.log 2>&1
cat $OUT/kbench_fused.log
