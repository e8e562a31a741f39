#!/bin/bash
# 8-GPU (or N) visit, short: headline c2 line + c4s strong scaling (parity asserted inside bench.py)
N=${1:-8}
mkdir -p gpurun_out
run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N "${@:3}"; }
run 600 29511 --steps 10 --warmup 3 > gpurun_out/bench_n${N}_c2.json 2> gpurun_out/bench_n${N}_c2.err; echo "c2 rc=$?"; tail -2 gpurun_out/bench_n${N}_c2.err
SECONDS=0
run 900 29513 --steps 5 --warmup 3 --workload c4s --sharded-only > gpurun_out/bench_n${N}_c4s.json 2> gpurun_out/bench_n${N}_c4s.err; echo "c4s rc=$? wall ${SECONDS}s"; tail -2 gpurun_out/bench_n${N}_c4s.err
