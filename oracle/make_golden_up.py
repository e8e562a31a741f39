"""TEST INFRASTRUCTURE ONLY -- golden tree for the reference's DEFAULT situation: depth PNGs at the sensor's
native size, masks / intrinsics at the (larger) colour size, so the unmodified reference resizes every depth image
with cv2.resize (tools/projection_2d_to_3d.py:436, :563; configs/config.yaml:19-20 are 1296x968 vs 640x480 depth).

Writes tests/golden/tree_up/ (inputs only) and tests/golden/e2e_up.pth (what the reference CLIs wrote, run through
oracle/run_reference.py in the build container). Run: python oracle/make_golden_up.py"""
from __future__ import annotations

import os
import shutil
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "beyond-fixed-forms_b200"), HERE, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import run_reference as RR  # noqa: E402
from make_golden import _pack_result  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
QUERY = "office chair"


def main():
    import yaml
    from bff_b200 import synth
    tree = os.path.join(GOLDEN, "tree_up")
    shutil.rmtree(tree, ignore_errors=True)
    os.makedirs(tree)
    scenes = {}
    for i, (seed, npts, nfr) in enumerate([(15, 3000, 7), (16, 2600, 6)]):
        sp = synth.SceneSpec(n_points=npts, n_frames=nfr, width=162, height=121, depth_width=80, depth_height=60,
                             n_boxes=9, n_stage1=12, seed=seed, n_queries=1)
        scenes[f"scene{i + 1:04d}_00"] = synth.make_scene(sp)
    cfg_path = synth.write_tree(tree, scenes, [QUERY])
    cfg = yaml.safe_load(open(cfg_path))
    for k, v in list(cfg.items()):
        if isinstance(v, str) and v.startswith(tree):
            cfg[k] = "." + v[len(tree):]
    yaml.safe_dump(cfg, open(cfg_path, "w"))
    expected = {}
    RR.run_cli("projection_2d_to_3d.py", "config.yaml", QUERY, tree)
    RR.run_cli("refinement.py", "config.yaml", QUERY, tree)
    for stage, d in (("mask_3d", cfg["mask_3d_dir"]), ("final", cfg["final_output_dir"])):
        qd = os.path.join(tree, d, QUERY)
        for fn in sorted(os.listdir(qd)):
            r = torch.load(os.path.join(qd, fn), map_location="cpu", weights_only=False)
            expected[("config.yaml", QUERY, stage, fn[:-4])] = _pack_result(r)
    torch.save(expected, os.path.join(GOLDEN, "e2e_up.pth"))
    for d in ("output/mask_3d", "output/final_output", "checkpoints"):
        shutil.rmtree(os.path.join(tree, d), ignore_errors=True)
    for k, v in expected.items():
        print(k, v["ins_shape"], len(v["final_class"]))


if __name__ == "__main__":
    main()
