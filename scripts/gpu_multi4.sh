#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n${N}.txt 2>&1
run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N "${@:3}"; }
SECONDS=0
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,GRAPH run 900 29513 --steps 5 --warmup 3 --workload c4s --sharded-only > gpurun_out/bench_n${N}_c4s.json 2> gpurun_out/bench_n${N}_c4s.err; echo "c4s rc=$? wall ${SECONDS}s"
grep -E "via|NVLS|Channel 00|Connected|nvls" gpurun_out/bench_n${N}_c4s.err | head -12
BFF_PEER_ROWS=0 run 900 29514 --steps 5 --warmup 3 --workload c4s --sharded-only > gpurun_out/bench_n${N}_c4s_nccl.json 2> gpurun_out/bench_n${N}_c4s_nccl.err; echo "c4s (nccl exchange) rc=$?"
