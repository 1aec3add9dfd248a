#!/bin/bash
# r2p: multi-GPU confirmation of the final tree (one-call link default) on an 8-GPU box: 2-rank NCCL counter test,
# weak-scaling bench lines at 2 / 4 / 8 GPUs, the reference arm under torchrun, the strong-scaling sweep at 8 GPUs.
TAG=${1:-r2p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.max.sm,memory.total --format=csv > $OUT/gpus.csv 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q -p no:cacheprovider -s > $OUT/pytest_dist.log 2>&1
echo "pytest dist exit $?" | tee -a $OUT/pytest_dist.log; tail -3 $OUT/pytest_dist.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29600
for n in 2 4 8; do
  P=$((P+1))
  timeout 600 $TR --nproc-per-node $n --master-port $P bench.py --gpus $n --steps 10 --warmup 3 > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err
  echo "bench n=$n exit $?"; cut -c1-330 $OUT/bench_n$n.json
done
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err; cut -c1-200 $OUT/bench_n1.json
P=$((P+1))
timeout 600 $TR --nproc-per-node 8 --master-port $P bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > $OUT/bench_reference_n8.json 2> $OUT/bench_reference_n8.err
echo "reference arm n=8 exit $?"; cut -c1-300 $OUT/bench_reference_n8.json
P=$((P+1))
timeout 900 $TR --nproc-per-node 8 --master-port $P bench.py --gpus 8 --sweep 1024,4096,16384,65536,262144,1048576 --sweep-out $OUT/sweep_n8.json > $OUT/sweep_n8.line 2> $OUT/sweep_n8.err
echo "sweep n=8 exit $?"; cut -c1-600 $OUT/sweep_n8.line
ls -la $OUT
