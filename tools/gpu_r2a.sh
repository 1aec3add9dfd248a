#!/bin/bash
# Round-2 first GPU pass: parity tests on the new tree, bench (both links), configs[4] sweep at N=1,
# per-waveform kbench, ncu full captures of the PCM/FM and multi-h chains (final round-1 kernels).
TAG=${1:-r2a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > $OUT/gpu.csv 2>&1
nproc > $OUT/host_cores.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/host_cores.txt
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -15 $OUT/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
echo "bench exit $?"; cat $OUT/bench.json
timeout 900 python bench.py --fused --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_fused.json 2> $OUT/bench_fused.err
echo "bench fused exit $?"; cat $OUT/bench_fused.json
timeout 1200 python bench.py --sweep 1024,4096,16384,65536,262144,1048576 --sweep-cpu 4096 --sweep-out $OUT/sweep_n1.json > $OUT/sweep_n1.line 2> $OUT/sweep_n1.err
echo "sweep exit $?"; tail -c 1500 $OUT/sweep_n1.line
for wf in pcmfm multih soqpsk-tg soqpsk-b; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf >> $OUT/kbench_waveforms.log 2>&1
done
cat $OUT/kbench_waveforms.log
for wf in pcmfm multih; do
  CMD="python tools/kbench.py --frames 32768 --reps 1 --waveform $wf"
  $CMD > $OUT/ncu_plain_$wf.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'viterbi_cpm|mf_bank' -s 4 -c 2 -o $OUT/prof_$wf $CMD > $OUT/ncu_full_$wf.log 2>&1
  echo "ncu $wf exit $?"
done
ls -la $OUT
