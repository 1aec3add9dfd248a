#!/bin/bash
# r2c: full GPU test pass, per-kernel timings of the three chains, ncu full captures of the new kernels
TAG=${1:-r2c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -30
for ex in '{"fused": true, "fused_tune": 1}' '{"fused": true, "fused_tune": 2}' '{}'; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 5 --extra "$ex" >> $OUT/kbench.log 2>&1
done
for wf in pcmfm multih; do
  for ex in '{}' '{"tune": 4}' '{"depth": 32}'; do
    timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf --extra "$ex" >> $OUT/kbench.log 2>&1
  done
done
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform soqpsk-b >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
# ncu: full captures (one launch each) of the fused SOQPSK kernel and the two fused detectors
CMD="python tools/kbench.py --frames 32768 --reps 1 --extra {\"fused\":true}"
$CMD > $OUT/ncu_plain_fused.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'cpm_modulate_bank_warp' -s 2 -c 1 -o $OUT/prof_fusedwarp $CMD > $OUT/ncu_fusedwarp.log 2>&1
echo "ncu fused exit $?"
for wf in pcmfm multih; do
  CMD="python tools/kbench.py --frames 32768 --reps 1 --waveform $wf"
  $CMD > $OUT/ncu_plain_$wf.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'cpm_detect' -s 2 -c 1 -o $OUT/prof_$wf $CMD > $OUT/ncu_$wf.log 2>&1
  echo "ncu $wf exit $?"
done
ls -la $OUT
