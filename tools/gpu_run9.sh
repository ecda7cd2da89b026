#!/bin/bash
set -u
mkdir -p gpurun_out
rm -f gpurun_out/tune.log
for lib in V4 S2P1 S1P2 S2P2; do BZ_B200_LIB=$PWD/scratch/lib_$lib.so timeout 300 python tools/tune_wave.py exact persistent 2>&1 | tee -a gpurun_out/tune.log; done
timeout 300 python tools/tune_wave.py exact persistent 21 22 23 24 2>&1 | tee -a gpurun_out/tune.log
timeout 300 python tools/tune_wave.py fast 22 24 2>&1 | tee -a gpurun_out/tune.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
python tools/tune_wave.py exact 22 > gpurun_out/plain_t2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_wave_eval -s 70 -c 1 -o gpurun_out/prof_eval -f python tools/tune_wave.py exact 22 > gpurun_out/ncu_t2.log 2>&1
tail -2 gpurun_out/ncu_t2.log
