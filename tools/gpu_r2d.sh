#!/bin/bash
# r2d: full GPU test pass on the tree with the fused link as default, bench (fused default / --unfused), sweep N=1, kbench
TAG=${1:-r2d}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > $OUT/gpu.csv 2>&1
nproc > $OUT/host_cores.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/host_cores.txt
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -8
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
echo "bench exit $?"; cat $OUT/bench.json
timeout 900 python bench.py --unfused --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_unfused.json 2> $OUT/bench_unfused.err
echo "bench unfused exit $?"; cat $OUT/bench_unfused.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err
echo "bench ref exit $?"; cat $OUT/bench_reference.json
timeout 1200 python bench.py --sweep 1024,4096,16384,65536,262144,1048576 --sweep-cpu 4096 --sweep-out $OUT/sweep_n1.json > $OUT/sweep_n1.line 2> $OUT/sweep_n1.err
echo "sweep exit $?"
for wf in soqpsk-tg pcmfm multih soqpsk-a soqpsk-mil; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf >> $OUT/kbench.log 2>&1
done
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform multih --extra '{"tune": 1}' >> $OUT/kbench.log 2>&1
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform pcmfm --extra '{"tune": 1}' >> $OUT/kbench.log 2>&1
cat $OUT/kbench.log
# launch list of the bench command (only after the identical command exited 0 without ncu)
NCU_CMD="python bench.py --frames 32768 --steps 2 --warmup 3 --no-cpu-baseline"
$NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $NCU_CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$NCU_CMD > $OUT/ncu_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_modulate|mf_bank|viterbi|encode_frames' -s 9 -c 3 -o $OUT/prof $NCU_CMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
