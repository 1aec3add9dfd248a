#!/bin/bash
TAG=${1:-r2t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
for i in 1 2; do timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform soqpsk-tg >> $OUT/kbench.log 2>&1; done
timeout 300 python tools/kbench.py --frames 65536 --reps 5 --waveform soqpsk-tg --unfused >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; cut -c1-330 $OUT/bench.json
