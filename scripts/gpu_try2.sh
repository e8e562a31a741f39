#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_fuzz_gpu.py -q -x > gpurun_out/pytest_lift.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_lift.log
timeout 600 python scripts/lift_variants.py c2 > gpurun_out/lift_variants.txt 2>&1; echo "variants rc=$?"
grep -E "fpc=  (25|30|38)" gpurun_out/lift_variants.txt
