#!/bin/bash
# what the round end runs: pytest -x -q -m gpu, smoke, bench
TAG=${1:-r3g}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1800 python -m pytest tests/ -x -q -m gpu > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $OUT/smoke.log
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cut -c1-260 $OUT/bench.json
