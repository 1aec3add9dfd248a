#!/bin/bash
TAG=${1:-r2q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_examples.py -m gpu -q -p no:cacheprovider -s > $OUT/pytest_examples.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_examples.log; tail -30 $OUT/pytest_examples.log
python examples/irig_comparison.py --bits 65536 --batch 16 > $OUT/irig_comparison.json 2> $OUT/irig.err; cat $OUT/irig_comparison.json
python examples/soqpsk_detection.py --frames 64 > $OUT/soqpsk_detection.json 2> $OUT/det.err; cat $OUT/soqpsk_detection.json
