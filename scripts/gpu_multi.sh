#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): headline scene-parallel c2 line (+ frame-sharded secondary), then the
# frame-sharded c3 (queries split over the ranks) and c4 (2M points, frames/rank fixed) workloads.
N=${1:-2}
mkdir -p gpurun_out
run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N "${@:3}"; }
run 600 29511 --steps 10 --warmup 3 > gpurun_out/bench_n${N}_c2.json 2> gpurun_out/bench_n${N}_c2.err; echo "c2 rc=$?"; cat gpurun_out/bench_n${N}_c2.json
run 600 29512 --steps 5 --warmup 3 --workload c3 > gpurun_out/bench_n${N}_c3.json 2> gpurun_out/bench_n${N}_c3.err; echo "c3 rc=$?"; cat gpurun_out/bench_n${N}_c3.json
run 900 29513 --steps 5 --warmup 3 --workload c4 > gpurun_out/bench_n${N}_c4.json 2> gpurun_out/bench_n${N}_c4.err; echo "c4 rc=$?"; cat gpurun_out/bench_n${N}_c4.json
tail -n 5 gpurun_out/bench_n${N}_c*.err
