#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'components_kernel|intersect_umma_kernel' -c 2 -o gpurun_out/comp_umma -f python bench.py --steps 1 --warmup 3 --profile > gpurun_out/ncu_comp.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
