"""GPU, >= 2 devices: ONE scene sharded over the ranks (frame-sharded lifting with the rows written in column slices, the
grouped NCCL row exchange + vote all-reduce of the library's own communicator, point-sharded device-resident pipelines
with int32 all-reduces inside) must give, bit for bit, the final masks of the same scene on one GPU. bench.py does the
comparison itself on rank 0 (frame_sharded_run raises on a mismatch) and reports it in its JSON line."""
import json
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _run(world, workload, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", str(world), "--workload", workload, "--steps", "3",
           "--warmup", "3", "--sharded-only"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-4000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    return line


@pytest.mark.parametrize("workload", ["tiny", "tiny3", "c1"])
def test_sharded_scene_equals_single_gpu(cuda_dev, workload):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    line = _run(world, workload, 29581 + ["tiny", "tiny3", "c1"].index(workload))
    fs = line["frame_sharded"]
    assert fs["final_masks_identical_to_single_gpu"] is True
    assert fs["scaling"] == "strong" and line["n_gpus"] == world
    assert fs["lifted_masks_total"] > 0
