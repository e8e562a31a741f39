#!/bin/bash
# Quick GPU visit while developing: all GPU tests, then bench lines + the launch list.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
if [ "$1" != "notest" ]; then timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; fi; echo "pytest gpu rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dev.json 2> gpurun_out/bench_dev.err; echo "bench dev rc=$?"
cat gpurun_out/bench_dev.json; tail -5 gpurun_out/bench_dev.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file gpurun_out/launches_dev.csv python bench.py --steps 2 --warmup 3 --profile > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
for wl in c3 c4 c4s; do
  SECONDS=0; timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl rc=$? wall ${SECONDS}s"
  tail -3 gpurun_out/bench_$wl.err
done
