#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_n1.json"))
for k in ("value","ms_per_step","flush_ms","wall_ms_per_step_with_flush","frame_crc32","gpu_launches","clocks"): print(k, d[k])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"]); print("moving", d["e2e_moving_camera"]); print("fast", d.get("fast")); print("cpu", d.get("cpu_baseline"))
r=d["roofline"]; print("roofline", r["achieved"], r["peak"], r["frac"], r["peak_nominal"], r["frac_of_nominal"], r["kernel_ms"])
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-600 gpurun_out/bench_ref.json
