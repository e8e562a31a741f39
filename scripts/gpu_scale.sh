#!/bin/bash
# Exactly what the driver launches for the N-GPU line (default workload), plus the reference arm under torchrun.
N=${1:-2}
mkdir -p gpurun_out
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n${N}.json 2> gpurun_out/scale_n${N}.err; echo "scale n$N rc=$? wall ${SECONDS}s"; tail -2 gpurun_out/scale_n${N}.err
