#!/bin/bash
# One gpurun call: ncu full capture of the four stage kernels of one waveform's chain (developer tool).
#   bash tools/ncu_waveform.sh <tag> <waveform> [frames]
TAG=${1:-wf}; WF=${2:-pcmfm}; FR=${3:-16384}
OUT=gpurun_out/$TAG; mkdir -p $OUT
CMD="python tools/kbench.py --frames $FR --reps 2 --waveform $WF"
timeout 600 $CMD > $OUT/plain_$WF.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain_$WF.log; exit 1; }
tail -1 $OUT/plain_$WF.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_modulate|mf_bank|viterbi|encode_frames' -s 8 -c 4 \
  -o $OUT/prof_$WF $CMD > $OUT/ncu_$WF.log 2>&1
echo "ncu exit $?"
