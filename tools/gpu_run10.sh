#!/bin/bash
set -u
mkdir -p gpurun_out
rm -f gpurun_out/tune.log
timeout 300 python tools/tune_wave.py exact persistent 22 23 24 2>&1 | tee -a gpurun_out/tune.log
timeout 300 python tools/tune_wave.py fast persistent 22 24 2>&1 | tee -a gpurun_out/tune.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
python tools/tune_wave.py exact persistent > gpurun_out/plain_t2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_warp_trace -s 2 -c 1 -o gpurun_out/prof_persist -f python tools/tune_wave.py exact persistent > gpurun_out/ncu_t2.log 2>&1
tail -2 gpurun_out/ncu_t2.log
python tools/tune_wave.py exact 22 > gpurun_out/plain_t3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 140 -c 70 --csv --log-file gpurun_out/launches_wave22.csv python tools/tune_wave.py exact 22 > gpurun_out/ncu_t3.log 2>&1
