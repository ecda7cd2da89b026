#!/bin/bash
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/bench_n2.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n2.json").read().strip().splitlines()[-1])
for k in ("value","ms_per_step","flush_ms","frame_crc32","gpu_launches"): print(k, d[k])
print("e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"]); print("moving", d["e2e_moving_camera"]["value"], d["e2e_moving_camera"]["plan_ms"])
print("roofline kernel ms", d["roofline"]["kernel_ms"], d["roofline"]["peak"])
PY
