#!/bin/bash
# one gpurun call: GPU tests, bench in both builds, the fast-mode report, ncu launch list + one full capture of the fast trace kernel
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_exact.json 2> gpurun_out/bench_exact.err; echo "bench exact rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --mode fast > gpurun_out/bench_fast.json 2> gpurun_out/bench_fast.err; echo "bench fast rc=$?"
cat gpurun_out/bench_exact.json gpurun_out/bench_fast.json | cut -c1-400
timeout 600 python tools/fast_report.py > gpurun_out/fast_report.jsonl 2> gpurun_out/fast_report.err; echo "fast report rc=$?"
cat gpurun_out/fast_report.jsonl
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --mode fast > gpurun_out/plain_fast.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_fast.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --mode fast > gpurun_out/ncu_l_fast.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --mode fast > gpurun_out/plain_fast2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_warp_trace -s 1 -c 1 -o gpurun_out/prof_fast -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --mode fast > gpurun_out/ncu_f_fast.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_exact.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_warp_trace -s 1 -c 1 -o gpurun_out/prof_exact -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_f_exact.log 2>&1
ls -la gpurun_out
