#!/bin/bash
set -u
mkdir -p gpurun_out
rm -f gpurun_out/tune.log
python scratch/time_r01.py 2>&1 | tee -a gpurun_out/tune.log
for lib in PLAIN V4 BOTH; do BZ_B200_LIB=$PWD/scratch/lib_$lib.so timeout 300 python tools/tune_wave.py exact persistent 2>&1 | tee -a gpurun_out/tune.log; done
timeout 300 python tools/tune_wave.py exact persistent 2>&1 | tee -a gpurun_out/tune.log
timeout 1500 python -m pytest tests/test_gpu_wide.py -m gpu -x -q -s 2>&1 | tail -30
