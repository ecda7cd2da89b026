#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for pipe in persistent wavefront; do for mode in exact fast; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --mode $mode --pipeline $pipe > gpurun_out/bench_${mode}_${pipe}.json 2> gpurun_out/bench_${mode}_${pipe}.err; echo "bench $mode $pipe rc=$?"
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_${mode}_${pipe}.json")); print("$mode $pipe", d["ms_per_step"], d["value"], d["roofline"]["kernel_ms"], d["e2e"]["value"], d["work_per_step"])
except Exception as e: print("no json", e); print(open("gpurun_out/bench_${mode}_${pipe}.err").read()[-2000:])
PY
done; done
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline wavefront > gpurun_out/plain_wave.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_wave.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline wavefront > gpurun_out/ncu_l_wave.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline wavefront > gpurun_out/plain_wave2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_wave -s 73 -c 6 -o gpurun_out/prof_wave -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline wavefront > gpurun_out/ncu_f_wave.log 2>&1
ls -la gpurun_out | tail -20
