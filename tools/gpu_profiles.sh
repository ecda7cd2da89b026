#!/bin/bash
# one gpurun call: the bench lines and ncu captures that go to profiles/ (r02)
set -u
O=gpurun_out/r02
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit,driver_version --format=csv > $O/gpu.txt 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_cfg4_n1.json 2> $O/bench_cfg4_n1.err; echo "cfg4 rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --pipeline persistent --no-cpu-baseline > $O/bench_cfg4_persistent.json 2> $O/err.txt; echo "persistent rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --mode fast --no-cpu-baseline > $O/bench_cfg4_fast.json 2>> $O/err.txt; echo "fast rc=$?"
for w in cfg1_sphere_640x480 cfg2_ellipsoid64x32_1080p cfg3_robot_1080p; do
  timeout 600 python bench.py --steps 20 --warmup 5 --workload $w --cpu-seconds 20 > $O/bench_$w.json 2>> $O/err.txt; echo "$w rc=$?"
done
timeout 600 python bench.py --steps 10 --warmup 3 --workload illum_ellipsoid64x32_4m > $O/bench_illum.json 2>> $O/err.txt; echo "illum rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_cfg4_reference_arm.json 2>> $O/err.txt; echo "ref rc=$?"
timeout 900 python tools/fast_report.py > $O/fast_report.jsonl 2>> $O/err.txt; echo "fast report rc=$?"
# launch lists (one pass of gpu__time_duration) and full captures, each after the same command ran without ncu
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/launches_wavefront.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu1.log 2>&1
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline persistent > $O/plain2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_persistent.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline persistent > $O/ncu2.log 2>&1
# DRAM bytes of every launch of two frames of the default pipeline (profiles/traffic.json)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"bz_" -s 28 -c 28 --csv --log-file $O/traffic_frame.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu7.log 2>&1
# the 12 wavefront launches of one frame (2 chunks x 2 crossings x 3 kernels)
ncu --set full --clock-control none --import-source on -k regex:bz_wave -s 12 -c 12 -o $O/prof_wave -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"bz_warp_trace|bz_sky" -s 2 -c 2 -o $O/prof_persistent -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --pipeline persistent > $O/ncu5.log 2>&1
python tools/tune_wave.py fast persistent > $O/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bz_warp_trace -s 2 -c 1 -o $O/prof_fast -f python tools/tune_wave.py fast persistent > $O/ncu3.log 2>&1
cat $O/err.txt | tail -5
ls -la $O
