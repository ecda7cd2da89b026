#!/bin/bash
# bench.py at N GPUs under torchrun (N = $1), as the driver launches it
set -u
N=$1
mkdir -p gpurun_out/r02
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02/bench_cfg4_n$N.json 2> gpurun_out/r02/bench_n$N.err; echo "bench n$N rc=$?"; tail -3 gpurun_out/r02/bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02/bench_cfg4_n$N.json").read().strip().splitlines()[-1])
print("N=$N value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "moving", d["e2e_moving_camera"]["value"], "crc", d["frame_crc32"], "kernel_ms", d["roofline"]["kernel_ms"])
PY
