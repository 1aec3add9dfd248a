#!/bin/bash
TAG=${1:-r2w}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for wf in pcmfm multih; do
  CMD="python tools/kbench.py --frames 32768 --reps 2 --waveform $wf"
  timeout 300 $CMD > $OUT/plain_$wf.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_detect' -s 2 -c 1 -o $OUT/prof_$wf $CMD > $OUT/ncu_$wf.log 2>&1
  echo "ncu $wf exit $?"
done
