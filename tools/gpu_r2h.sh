#!/bin/bash
# r2h: MIL bank instances + full suite + kbench of every waveform on the final tree
TAG=${1:-r2h}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -6
for wf in soqpsk-tg pcmfm multih soqpsk-a soqpsk-mil soqpsk-b; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf >> $OUT/kbench.log 2>&1
done
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform soqpsk-tg --unfused >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
