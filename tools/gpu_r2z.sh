#!/bin/bash
# r2z: evidence pass on the final tree: full GPU suite, smoke, bench (fused default / --unfused / --impl reference),
# ncu launch list + full capture of the bench command, full captures of the PCM/FM and ARTM fused detectors.
TAG=${1:-r2z}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > $OUT/gpu.csv 2>&1
nproc > $OUT/host_cores.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/host_cores.txt
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -s > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/pytest_gpu.log
grep -v "float32 flips" $OUT/pytest_gpu.log | tail -4
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cut -c1-400 $OUT/bench.json
timeout 900 python bench.py --unfused --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_unfused.json 2> $OUT/bench_unfused.err; echo "bench unfused exit $?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench ref exit $?"; cut -c1-300 $OUT/bench_reference.json
NCU_CMD="python bench.py --frames 32768 --steps 2 --warmup 3 --no-cpu-baseline"
$NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $NCU_CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$NCU_CMD > $OUT/ncu_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_modulate|mf_bank|viterbi|encode_frames' -s 9 -c 3 -o $OUT/prof $NCU_CMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
for wf in pcmfm multih; do
  CMD="python tools/kbench.py --frames 32768 --reps 2 --waveform $wf"
  timeout 300 $CMD > $OUT/plain_$wf.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_detect' -s 2 -c 1 -o $OUT/prof_$wf $CMD > $OUT/ncu_$wf.log 2>&1
  echo "ncu $wf exit $?"
done
ls -la $OUT
