#!/bin/bash
# r2f: tree with diagnostics (diag.cu), the one-call link (link.cu) and the static-scratch detector (noalloc.cu):
# new tests first (fast feedback), test_zz_no_alloc alone in a fresh process, then the full GPU suite, smoke, bench.
TAG=${1:-r2g}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > $OUT/gpu.csv 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider -s > $OUT/pytest_noalloc_alone.log 2>&1
echo "noalloc alone exit $?" | tee -a $OUT/pytest_noalloc_alone.log; tail -5 $OUT/pytest_noalloc_alone.log
timeout 600 python -m pytest tests/test_gpu_link.py tests/test_gpu_diag.py -m gpu -q -p no:cacheprovider -s > $OUT/pytest_new.log 2>&1
echo "new tests exit $?" | tee -a $OUT/pytest_new.log; tail -15 $OUT/pytest_new.log
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -12
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log
nproc > $OUT/host_cores.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/host_cores.txt
timeout 300 python oracle/ref_bench.py --frames-single 8 --frames-pool-per-core 8 > $OUT/ref_bench.json 2> $OUT/ref_bench.err; echo "ref_bench exit $?"; cat $OUT/ref_bench.json
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
echo "bench exit $?"; cat $OUT/bench.json
timeout 600 python bench.py --sweep 1024,4096,16384,65536 --sweep-cpu 0 --sweep-out $OUT/sweep_n1.json > $OUT/sweep_n1.line 2> $OUT/sweep_n1.err
echo "sweep exit $?"; cat $OUT/sweep_n1.line | head -c 1500
ls -la $OUT
