#!/bin/bash
# round 2, last GPU call: the allocation-free entry point's test alone (fresh process), then the whole GPU suite (it is
# collected last there), smoke(), and a short default bench line without the CPU leg
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/pytest_zz_alone.log
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 200 python bench.py --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"; tail -2 gpurun_out/bench_c2.err
python - <<'PY'
import json
try:
    d = json.loads([l for l in open("gpurun_out/bench_c2.json") if l.startswith("{")][-1])
    print(d["ms_per_step"], d["stages_ms"], d["roofline"]["ms_per_launch"], d["roofline"]["frac"], d["e2e"]["ms_per_step"])
except Exception as e:
    print("bench line unreadable", e)
PY
