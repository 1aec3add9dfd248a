#!/bin/bash
# r3c: BER waterfalls of the PCM/FM and ARTM links with the reachable-class detectors (one GPU, 65536-bit frames)
TAG=${1:-r3c}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python examples/soqpsk_ber_waterfall.py --waveform pcmfm --ebn0 0 1 2 3 4 5 6 7 8 9 10 11 12 --n-bits 65536 --batch 4096 --max-bits 2e10 --out $OUT/waterfall_pcmfm.json > /dev/null 2> $OUT/waterfall_pcmfm.txt; echo "pcmfm exit $?"; tail -15 $OUT/waterfall_pcmfm.txt
timeout 600 python examples/soqpsk_ber_waterfall.py --waveform multih --ebn0 0 2 4 6 8 9 10 11 12 13 14 --n-bits 65536 --batch 4096 --max-bits 2e10 --out $OUT/waterfall_multih.json > /dev/null 2> $OUT/waterfall_multih.txt; echo "multih exit $?"; tail -13 $OUT/waterfall_multih.txt
