#!/bin/bash
set -u
O=gpurun_out/r02
mkdir -p $O
timeout 600 python bench.py --steps 10 --warmup 3 --pipeline wavefront --no-cpu-baseline > $O/bench_cfg4_wavefront.json 2> $O/err2.txt; echo "wave rc=$?"
python -c "
import json; d=json.load(open('$O/bench_cfg4_wavefront.json')); print('wavefront', d['value'], d['ms_per_step'], d['e2e']['value'], d['frame_crc32'], d['clocks'])"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_check.json 2>> $O/err2.txt; python -c "
import json; d=json.load(open('$O/bench_check.json')); print('default', d['value'], d['ms_per_step'], d['clocks'], d['e2e_moving_camera'])"
BZ_TUNE_WORKLOAD=cfg3_robot_1080p python tools/tune_wave.py exact persistent > $O/plain6.log 2>&1 &&
BZ_TUNE_WORKLOAD=cfg3_robot_1080p ncu --set full --clock-control none --import-source on -k regex:bz_warp_trace -s 2 -c 1 -o $O/prof_cfg3 -f python tools/tune_wave.py exact persistent > $O/ncu6.log 2>&1
cat $O/plain6.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
