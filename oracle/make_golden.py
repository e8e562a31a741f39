"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ from the UNMODIFIED reference.

Run in the build container (where /root/reference exists):  python oracle/make_golden.py
Reference commit: 1d3aff0f. Everything written here is consumed by tests/ only.

Two kinds of fixtures:
  functions.npz   inputs + outputs of the reference FUNCTIONS imported from
                  tools/projection_2d_to_3d.py (:37-92, :149-166, :250-301), tools/refinement.py
                  (:26-39, :69-90) and tools/utils/rle_encode_decode.py (:10-61)
  tree/ + e2e.pth a small synthetic input tree in the reference's on-disk formats and the .pth
                  files the reference CLIs (projection_2d_to_3d.py, refinement.py run through
                  runpy, oracle/run_reference.py) wrote for it, masks bit-packed
"""
from __future__ import annotations

import os
import shutil
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "beyond-fixed-forms_b200"), HERE, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import run_reference as RR  # noqa: E402
from helpers import random_scene  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
QUERIES = ["office chair", "table"]


def function_goldens():
    P, R, E = RR.import_reference()
    out = {}
    sc = random_scene(77, n=2000, f=4, h=60, w=80, m_per_frame=(1, 2, 3))
    sc["xyz"][:6] = [[0, 0, 0], [1e30, 0, 0], [-1e9, 1e9, 3], [6, 5, 2.8], [1e-30, 1e-30, 1e-30], [3, 2, 1]]
    sc["xyz"][6] = sc["poses"][0][:3, 3]                     # on the camera centre of frame 0 (z = 0)
    sc["poses"][2][:] = -np.inf                              # a broken ScanNet pose file
    for k in ("xyz", "K", "poses", "depth", "mask_off"):
        out["in_" + k] = sc[k]
    out["in_masks"] = np.packbits(sc["masks"].reshape(sc["masks"].shape[0], -1), axis=1)
    pcd = np.concatenate([sc["xyz"], torch.ones([sc["n"], 1])], axis=1).T          # :385-390
    pix_all, vis_all, cam_all, rows_all, pinv_all = [], [], [], [], []
    for f in range(sc["f"]):
        with np.errstate(all="ignore"):
            pinv = np.linalg.inv(sc["poses"][f])                                   # :425
            scene_pts = (pinv @ pcd).T[:, :3]
            pix = P.compute_projected_pts_tensor(scene_pts, sc["K"])               # :37
            vis = P.compute_visibility_mask_tensor(scene_pts, pix, sc["depth"][f], 0.08)   # :51
        m0, m1 = sc["mask_off"][f], sc["mask_off"][f + 1]
        rows = P.compute_visible_masked_pts_tensor(scene_pts, pix, vis, sc["masks"][m0:m1].astype(np.float32))  # :73
        pinv_all.append(pinv); cam_all.append(scene_pts); pix_all.append(pix); vis_all.append(vis); rows_all.append(rows)
    out["pinv"], out["cam"], out["pix"], out["vis"] = (np.stack(x) for x in (pinv_all, cam_all, pix_all, vis_all))
    rows = np.concatenate(rows_all, axis=0)
    out["rows"] = np.packbits(rows, axis=1)
    # IoU (:149-166) incl. two empty rows, and the stage-to-stage form (refinement.py:69-90)
    rng = np.random.default_rng(3)
    a = rng.random((23, 1501)) < rng.uniform(0, 0.3, size=(23, 1))
    a[4] = False
    a[11] = False
    b = rng.random((7, 1501)) < 0.2
    out["iou_a"], out["iou_b"] = np.packbits(a, axis=1), np.packbits(b, axis=1)
    out["iou_aa"] = P.calculate_iou(torch.from_numpy(a)).numpy()
    out["iou_ab_T"] = R.calculate_iou_between_stages(torch.from_numpy(a).to(torch.uint8), torch.from_numpy(b)).numpy()
    # components (:250-274) and overlap resolution (:277-301) on the thresholded IoU graph
    adj = torch.from_numpy(out["iou_aa"]) > 0.12
    out["adj"] = adj.numpy()
    comps = P.find_unconnected_subgraphs_tensor(adj.float())
    out["components"] = np.array([",".join(map(str, c)) for c in comps])
    groups = [c for c in comps if len(c) >= 1]
    merged = torch.stack([torch.from_numpy(a[c]).any(0) for c in groups])
    out["overlap_in"] = np.packbits(merged.numpy(), axis=1)
    out["overlap_groups"] = np.array([",".join(map(str, c)) for c in groups])
    out["overlap_out"] = np.packbits(P.solve_overlapping(merged.clone(), groups).numpy(), axis=1)
    # RLE codec (rle_encode_decode.py:10-61, refinement.py:26-39) + a hostile hand-made RLE
    m = rng.random((5, 333)) < 0.4
    enc = E.rle_encode_batch(torch.from_numpy(m))
    out["rle_masks"] = m
    out["rle_counts"] = np.array([",".join(map(str, e["counts"].tolist())) for e in enc])
    dec = E.rle_decode_batch(enc)
    assert np.array_equal(dec.numpy().astype(bool), m)       # the reference's commented-out property (:124-128)
    hostile = {"length": 100, "counts": np.array([5, 10, 8, 30, 95, 50, 0, 3], np.int64)}
    out["rle_hostile_counts"] = hostile["counts"]
    out["rle_hostile_decoded"] = R.rle_decode(hostile)
    np.savez_compressed(os.path.join(GOLDEN, "functions.npz"), **out)
    print("functions.npz:", {k: getattr(v, "shape", None) for k, v in out.items()})


def _pack_result(d):
    """{ins, conf, final_class} as saved by the reference -> small plain-python/numpy record."""
    ins, conf = d["ins"], d["conf"]
    rec = {"final_class": list(d["final_class"])}
    if isinstance(ins, torch.Tensor):
        rec["ins_shape"] = tuple(ins.shape)
        rec["ins_dtype"] = str(ins.dtype)
        rec["ins_bits"] = np.packbits(ins.numpy().astype(bool), axis=1) if ins.numel() else np.zeros((0, 0), np.uint8)
    else:
        rec["ins_shape"], rec["ins_dtype"], rec["ins_bits"] = None, "list", None
    rec["conf"] = conf.numpy() if isinstance(conf, torch.Tensor) else list(conf)
    rec["conf_is_tensor"] = isinstance(conf, torch.Tensor)
    return rec


def tree_goldens():
    """Three scenes x two queries: a normal one, one where a query has no 2D masks at all in a scene
    (empty stage 2 -> 'other stage-1 masks' path), and one with the occurrence filter config."""
    import yaml
    from bff_b200 import synth
    tree = os.path.join(GOLDEN, "tree")
    shutil.rmtree(tree, ignore_errors=True)
    os.makedirs(tree)
    scenes = {}
    for i, (seed, npts, nfr) in enumerate([(5, 3000, 8), (6, 2500, 6), (9, 2000, 7)]):
        sp = synth.SceneSpec(n_points=npts, n_frames=nfr, width=160, height=120, n_boxes=9, n_stage1=12,
                             seed=seed, n_queries=2)
        scenes[f"scene{i + 1:04d}_00"] = synth.make_scene(sp)
    cfg_path = synth.write_tree(tree, scenes, QUERIES)
    # scene0003: query "table" saw nothing (file present, empty list) -> 'No 3d masks detected in backprojection'
    torch.save([], os.path.join(tree, "output", "mask_2d", "rle", "table", "scene0003_00.pth"))
    cfg = yaml.safe_load(open(cfg_path))
    for k, v in list(cfg.items()):
        if isinstance(v, str) and v.startswith(tree):
            cfg[k] = "." + v[len(tree):]                       # relative: tests run with cwd = a copy of the tree
    yaml.safe_dump(cfg, open(cfg_path, "w"))
    cfg_occ = dict(cfg, if_occurance_threshold=True, if_detected_ratio_threshold=False,
                   mask_3d_dir="./output_occ/mask_3d", final_output_dir="./output_occ/final_output")
    yaml.safe_dump(cfg_occ, open(os.path.join(tree, "config_occ.yaml"), "w"))
    expected = {}
    for cfg_name, c in (("config.yaml", cfg), ("config_occ.yaml", cfg_occ)):
        for q in QUERIES:
            RR.run_cli("projection_2d_to_3d.py", cfg_name, q, tree)
            RR.run_cli("refinement.py", cfg_name, q, tree)
            for stage, d in (("mask_3d", c["mask_3d_dir"]), ("final", c["final_output_dir"])):
                qd = os.path.join(tree, d, q)
                for fn in sorted(os.listdir(qd)):
                    r = torch.load(os.path.join(qd, fn), map_location="cpu", weights_only=False)
                    expected[(cfg_name, q, stage, fn[:-4])] = _pack_result(r)
    torch.save(expected, os.path.join(GOLDEN, "e2e.pth"))
    # keep only INPUTS in the committed tree
    for d in ("output/mask_3d", "output/final_output", "output_occ", "checkpoints"):
        shutil.rmtree(os.path.join(tree, d), ignore_errors=True)
    for k, v in expected.items():
        print(k, v["ins_shape"], len(v["final_class"]))


if __name__ == "__main__":
    os.makedirs(GOLDEN, exist_ok=True)
    function_goldens()
    tree_goldens()
