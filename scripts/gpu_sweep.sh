#!/bin/bash
# c5 (312-scene sweep, scene-parallel LPT) on N GPUs, launched like the driver launches bench.py
N=${1:-2}
mkdir -p gpurun_out
SECONDS=0
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --workload c5 > gpurun_out/sweep_n${N}.json 2> gpurun_out/sweep_n${N}.err; echo "sweep n$N rc=$? wall ${SECONDS}s"; tail -2 gpurun_out/sweep_n${N}.err; cut -c1-330 gpurun_out/sweep_n${N}.json
