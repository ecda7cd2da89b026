#!/bin/bash
# bench.py at N GPUs under torchrun (N = $1), as the driver launches it; $2 (optional): extra bench.py arguments, e.g. "--gather nccl"
set -u
N=$1
X=${2:-}
T=$(echo "$X" | tr -d ' -')
mkdir -p gpurun_out/r02
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 $X > gpurun_out/r02/bench_cfg4_n$N$T.json 2> gpurun_out/r02/bench_n$N$T.err; echo "bench n$N $X rc=$?"; tail -3 gpurun_out/r02/bench_n$N$T.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02/bench_cfg4_n$N$T.json").read().strip().splitlines()[-1])
print("N=$N $X value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "moving", d["e2e_moving_camera"]["value"], "crc", d["frame_crc32"], "kernel_ms", d["roofline"]["kernel_ms"], d["step_split_rank0"], d.get("gather"))
PY
