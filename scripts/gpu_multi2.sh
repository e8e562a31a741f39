#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): sharded-scene parity tests (+ the 2-rank CLI parity test at N=2), the headline bench
# line (c2: strong scaling + replicas), then the strong-scaling lines of c3 (20 queries) and c4s (2M points, 2000 frames).
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/gpu_n${N}.txt 2>&1
if [ "$2" != "notest" ]; then
timeout 900 python -m pytest tests/test_sharded_gpu.py -q -x > gpurun_out/pytest_sharded_n${N}.log 2>&1; echo "sharded pytest rc=$?" | tee -a gpurun_out/pytest_sharded_n${N}.log
tail -5 gpurun_out/pytest_sharded_n${N}.log
fi
run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N "${@:3}"; }
run 600 29511 --steps 10 --warmup 3 > gpurun_out/bench_n${N}_c2.json 2> gpurun_out/bench_n${N}_c2.err; echo "c2 rc=$?"; tail -3 gpurun_out/bench_n${N}_c2.err
for wl in c3 c4s; do
  SECONDS=0
  run 900 29513 --steps 5 --warmup 3 --workload $wl --sharded-only > gpurun_out/bench_n${N}_${wl}.json 2> gpurun_out/bench_n${N}_${wl}.err; echo "$wl rc=$? wall ${SECONDS}s"; tail -3 gpurun_out/bench_n${N}_${wl}.err
done
