#!/bin/bash
# fused CPM detectors + warp-autonomous fused SOQPSK kernel: parity tests + per-kernel timings
TAG=${1:-r2b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider -k "cpm or generic or fused or confidence" -s > $OUT/pytest_sel.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_sel.log
grep -v "float32 flips" $OUT/pytest_sel.log | tail -60
for ex in '{"fused": true, "fused_tune": 0}' '{"fused": true, "fused_tune": 1}' '{"fused": true, "fused_tune": 2}' '{}'; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 5 --extra "$ex" >> $OUT/kbench_fused.log 2>&1
done
cat $OUT/kbench_fused.log
for wf in pcmfm multih; do
  for ex in '{}' '{"tune": 2}' '{"tune": 4}' '{"depth": 8}' '{"depth": 32}' '{"unfused_detector": true}'; do
    timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf --extra "$ex" >> $OUT/kbench_cpm.log 2>&1
  done
done
cat $OUT/kbench_cpm.log
