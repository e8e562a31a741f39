"""In-process orchestration over classes: what evaluation/eval/run_evl.py:56-118 does with two fresh
Python processes per class -- tools/projection_2d_to_3d.py then tools/refinement.py -- done here for BATCHES of classes in
one process: per scene the cloud, the depth frames and the poses are loaded and uploaded once for the whole batch, one
lifting launch serves all its classes, the stage-1 masks are decoded once, and CLIP (when installed) is loaded once with a
cache of (query, label) similarities that lives across classes and scenes. The stage files and the orchestrator's own
checkpoint (process_checkpoint.txt, {class: {stage: true}}) are the reference's, so a run can be resumed by either tool;
the 2D segmentation and the evaluator stay the reference's and are not started here."""
from __future__ import annotations

import os
from typing import List, Sequence

import yaml

CHECKPOINT_FILE = "process_checkpoint.txt"               # run_evl.py:28


def read_checkpoint(path: str = CHECKPOINT_FILE) -> dict:
    if os.path.exists(path):
        with open(path) as fh:
            return yaml.safe_load(fh) or {}
    return {}


def write_checkpoint(ckpt: dict, path: str = CHECKPOINT_FILE):
    with open(path, "w") as fh:
        yaml.safe_dump(ckpt, fh)


def batches(items: Sequence[str], size: int) -> List[List[str]]:
    size = max(1, int(size))
    return [list(items[i:i + size]) for i in range(0, len(items), size)]


def run(config_path: str, classes: Sequence[str], batch: int = 20, stage2_main=None, refine_main=None) -> dict:
    """Both stages for `classes`, `batch` classes per pass over the scenes. Stages a class has already finished
    (process_checkpoint.txt) are skipped like run_evl.py does. Returns the checkpoint dict."""
    if stage2_main is None or refine_main is None:
        from . import refine, stage2
        stage2_main, refine_main = stage2_main or stage2.main_multi, refine_main or refine.main_multi
    ckpt = read_checkpoint()
    for cls in classes:
        ckpt.setdefault(cls, {})
    for group in batches(list(classes), batch):
        todo = [c for c in group if not ckpt[c].get("projection_2d_to_3d", False)]
        if todo:
            if stage2_main(["--config", config_path, "--cls"] + todo) != 0:
                raise RuntimeError(f"lifting failed for {todo}")
            for c in todo:
                ckpt[c]["projection_2d_to_3d"] = True
            write_checkpoint(ckpt)
        todo = [c for c in group if not ckpt[c].get("refinement", False)]
        if todo:
            if refine_main(["--config", config_path, "--cls"] + todo) != 0:
                raise RuntimeError(f"refinement failed for {todo}")
            for c in todo:
                ckpt[c]["refinement"] = True
            write_checkpoint(ckpt)
    return ckpt


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description="BeyondFF lifting + refinement for many classes, in-process (run_evl.py's stages 2 and 3)")
    ap.add_argument("--config", type=str, default="configs/config.yaml")
    ap.add_argument("--cls", type=str, nargs="+", required=True, help="classes (text queries), in processing order")
    ap.add_argument("--batch", type=int, default=20, help="classes that share one pass over the scenes")
    args = ap.parse_args(argv)
    run(args.config, args.cls, args.batch)
    return 0
