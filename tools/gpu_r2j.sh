#!/bin/bash
# r2j: run-time-tap bank instances (SOQPSK-B), issue roofline in the bench line
TAG=${1:-r2j}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -4
for wf in soqpsk-b soqpsk-mil soqpsk-tg; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf >> $OUT/kbench.log 2>&1
done
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform soqpsk-b --extra '{"detector": "PT"}' >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; python -c "
import json; b=json.load(open('$OUT/bench.json')); print(b['value'], b['roofline']['issue']); print({k:v['issue'] and round(v['issue']['frac'],3) for k,v in b['kernels'].items()})"
