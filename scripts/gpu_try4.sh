#!/bin/bash
mkdir -p gpurun_out
for wl in c3 c4 c4s; do
  SECONDS=0
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl rc=$? wall ${SECONDS}s"
  cat gpurun_out/bench_$wl.json; tail -3 gpurun_out/bench_$wl.err
done
