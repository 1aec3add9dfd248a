#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, then ncu (launch list + one full capture).
# Usage (from the repo root on the GPU box):  bash tools/gpu_check.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > $OUT/gpu.csv 2>&1
nproc > $OUT/host_cores.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $OUT/host_cores.txt

timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > $OUT/pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log

timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log; tail -2 $OUT/smoke.log

timeout 600 python bench.py --frames 4096 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/bench_small.json 2> $OUT/bench_small.err
echo "bench_small exit $?"; tail -c 600 $OUT/bench_small.json
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
echo "bench exit $?"; cat $OUT/bench.json
timeout 900 python bench.py --fused --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_fused.json 2> $OUT/bench_fused.err
echo "bench fused exit $?"; cat $OUT/bench_fused.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err
echo "bench ref exit $?"

# ncu: only after the identical command exited 0 without ncu
NCU_CMD="python bench.py --frames ${NCU_FRAMES:-32768} --steps 2 --warmup 3 --no-cpu-baseline"
$NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $NCU_CMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
$NCU_CMD > $OUT/ncu_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'cpm_modulate|mf_bank|viterbi|encode_frames' -s 12 -c 4 -o $OUT/prof $NCU_CMD > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT

# configs[1] / configs[2] of BASELINE.json (PCM/FM and ARTM multi-h, 65536 frames): per-kernel timings
for wf in pcmfm multih soqpsk-tg; do
  timeout 300 python tools/kbench.py --frames 65536 --reps 3 --waveform $wf >> $OUT/kbench_waveforms.log 2>&1
done
timeout 300 python tools/kbench.py --frames 65536 --reps 3 --fused >> $OUT/kbench_waveforms.log 2>&1
cat $OUT/kbench_waveforms.log
# launch list of the fused link (only after the identical command exited 0 without ncu)
FCMD="python bench.py --fused --frames ${NCU_FRAMES:-32768} --steps 2 --warmup 3 --no-cpu-baseline"
$FCMD > $OUT/ncu_plain_fused.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_fused.csv $FCMD > $OUT/ncu_launches_fused.log 2>&1
echo "ncu fused launches exit $?"
$FCMD > $OUT/ncu_plain_fused2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cpm_modulate -s 3 -c 1 -o $OUT/prof_fused $FCMD > $OUT/ncu_full_fused.log 2>&1
echo "ncu fused full exit $?"
