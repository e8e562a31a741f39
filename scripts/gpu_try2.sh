#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/lift_variants.py c2 > gpurun_out/lift_variants.txt 2>&1; echo "variants rc=$?"
cat gpurun_out/lift_variants.txt
BFF_LIB_PATH=$PWD/variants/c_pf_mb6.so timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:'lift_kernel' -c 1 -o gpurun_out/lift_pf_mb6 -f python bench.py --steps 1 --warmup 3 --profile > gpurun_out/ncu_lift.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
