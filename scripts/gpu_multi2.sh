#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): sharded-scene parity tests (+ the 2-rank CLI parity test at N=2), then the bench line.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/gpu_n${N}.txt 2>&1
timeout 900 python -m pytest tests/test_sharded_gpu.py -q -x > gpurun_out/pytest_sharded_n${N}.log 2>&1; echo "sharded pytest rc=$?" | tee -a gpurun_out/pytest_sharded_n${N}.log
tail -25 gpurun_out/pytest_sharded_n${N}.log
if [ "$N" = "2" ]; then
  timeout 600 python -m pytest tests/test_e2e_gpu.py -q -k two_rank > gpurun_out/pytest_two_rank.log 2>&1; echo "two-rank pytest rc=$?" | tee -a gpurun_out/pytest_two_rank.log
  tail -5 gpurun_out/pytest_two_rank.log
fi
run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N "${@:3}"; }
run 600 29511 --steps 10 --warmup 3 > gpurun_out/bench_n${N}_c2.json 2> gpurun_out/bench_n${N}_c2.err; echo "c2 rc=$?"; cat gpurun_out/bench_n${N}_c2.json; tail -15 gpurun_out/bench_n${N}_c2.err
