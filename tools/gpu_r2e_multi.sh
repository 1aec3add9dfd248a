#!/bin/bash
# r2e: multi-GPU pass on an 8-GPU box: 2-rank NCCL counter test on real kernels, weak-scaling bench lines,
# strong-scaling configs[4] sweep, and the BER waterfall to 1e-8 (BASELINE configs[3]).
TAG=${1:-r2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.max.sm,memory.total --format=csv > $OUT/gpus.csv 2>&1
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q -p no:cacheprovider -s > $OUT/pytest_dist.log 2>&1
echo "pytest dist exit $?" | tee -a $OUT/pytest_dist.log; tail -3 $OUT/pytest_dist.log
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
P=29500
for n in 2 4 8; do
  P=$((P+1))
  timeout 600 $TR --nproc-per-node $n --master-port $P bench.py --gpus $n --steps 10 --warmup 3 > $OUT/bench_n$n.json 2> $OUT/bench_n$n.err
  echo "bench n=$n exit $?"; cut -c1-400 $OUT/bench_n$n.json
done
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err
for n in 2 4 8; do
  P=$((P+1))
  timeout 900 $TR --nproc-per-node $n --master-port $P bench.py --gpus $n --sweep 1024,4096,16384,65536,262144,1048576 --sweep-out $OUT/sweep_n$n.json > $OUT/sweep_n$n.line 2> $OUT/sweep_n$n.err
  echo "sweep n=$n exit $?"
done
timeout 600 python bench.py --sweep 1024,4096,16384,65536,262144,1048576 --sweep-out $OUT/sweep_n1.json > $OUT/sweep_n1.line 2> $OUT/sweep_n1.err
# BER waterfall, 8 GPUs: the reference's frame length (floor from the frame-start decision), then longer frames
P=$((P+1))
timeout 600 $TR --nproc-per-node 8 --master-port $P examples/soqpsk_ber_waterfall.py --ebn0 0 1 2 3 4 5 6 7 8 9 10 11 12 13 14 15 16 --max-bits 2e10 --out $OUT/waterfall_pam_n4096.json > /dev/null 2> $OUT/waterfall_pam_n4096.txt
echo "waterfall 4096 exit $?"; cat $OUT/waterfall_pam_n4096.txt | tail -20
P=$((P+1))
timeout 900 $TR --nproc-per-node 8 --master-port $P examples/soqpsk_ber_waterfall.py --ebn0 0 1 2 3 4 5 6 7 8 9 10 11 12 13 14 15 16 17 --n-bits 1048576 --batch 256 --max-bits 1e11 --out $OUT/waterfall_pam_n1048576.json > /dev/null 2> $OUT/waterfall_pam_n1048576.txt
echo "waterfall 1M exit $?"; cat $OUT/waterfall_pam_n1048576.txt | tail -22
P=$((P+1))
timeout 900 $TR --nproc-per-node 8 --master-port $P examples/soqpsk_ber_waterfall.py --detector PT --ebn0 0 2 4 6 8 10 12 13 14 15 16 17 18 --n-bits 1048576 --batch 256 --max-bits 1e11 --out $OUT/waterfall_pt_n1048576.json > /dev/null 2> $OUT/waterfall_pt_n1048576.txt
echo "waterfall PT 1M exit $?"; cat $OUT/waterfall_pt_n1048576.txt | tail -16
ls -la $OUT
