#!/bin/bash
set -u
mkdir -p gpurun_out
for lib in scratch/lib_regs121.so scratch/lib_regs96.so; do
  BZ_B200_LIB=$PWD/$lib timeout 300 python tools/tune_wave.py exact persistent 17 18 19 20 22 2>&1 | tee -a gpurun_out/tune.log
done
BZ_B200_LIB=$PWD/scratch/lib_regs121.so timeout 300 python tools/tune_wave.py fast persistent 18 19 20 2>&1 | tee -a gpurun_out/tune.log
python tools/tune_wave.py exact 19 > gpurun_out/plain_t.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_wave -s 600 -c 6 -o gpurun_out/prof_wave19 -f python tools/tune_wave.py exact 19 > gpurun_out/ncu_t.log 2>&1
tail -3 gpurun_out/ncu_t.log
