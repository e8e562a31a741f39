#!/bin/bash
# round 2, late: AABB culling in the lift kernel (vs the sphere build in variants/), the c5 sweep, full GPU test suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python scripts/lift_variants.py c2 > gpurun_out/lift_variants_c2.txt 2>&1; echo "variants rc=$?"; grep "fpc=  30\|fpc=  25" gpurun_out/lift_variants_c2.txt
timeout 600 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; tail -3 gpurun_out/bench_c5.err; cut -c1-600 gpurun_out/bench_c5.json
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"; tail -2 gpurun_out/bench_c2.err
python - <<'PY'
import json
for f in ("gpurun_out/bench_c2.json",):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, d["ms_per_step"], d["stages_ms"], d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d["e2e"]["ms_per_step"])
    except Exception as e:
        print(f, "unreadable", e)
PY
