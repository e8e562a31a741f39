"""GPU: the reference's importable functions, served by libbff.so (bff_b200/reference_api.py, re-exported by
tools/projection_2d_to_3d.py and tools/refinement.py), called exactly like oracle/make_golden.py called the
UNMODIFIED reference, must reproduce tests/golden/functions.npz bit for bit."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.fixture(scope="module")
def G():
    return np.load(os.path.join(GOLDEN, "functions.npz"), allow_pickle=False)


def _bits(packed, n):
    return np.unpackbits(packed, axis=1)[:, :n].astype(bool)


def test_lifting_functions(cuda_dev, G):
    P = _load(os.path.join(ROOT, "tools", "projection_2d_to_3d.py"), "bff_tools_projection")
    depth, K = G["in_depth"], G["in_K"]
    f, h, w = depth.shape
    masks = _bits(G["in_masks"], h * w).reshape(-1, h, w)
    off = G["in_mask_off"]
    rows_g = _bits(G["rows"], G["in_xyz"].shape[0])
    for i in range(f):
        scene_pts = G["cam"][i]                                   # camera-space points as the reference passes them (:425-426)
        pix = P.compute_projected_pts_tensor(scene_pts, K)
        assert pix.dtype == np.int64 and pix.shape == (scene_pts.shape[0], 2)
        np.testing.assert_array_equal(pix, G["pix"][i])
        vis = P.compute_visibility_mask_tensor(scene_pts, pix, depth[i], 0.08)
        assert vis.dtype == bool
        np.testing.assert_array_equal(vis, G["vis"][i])
        rows = P.compute_visible_masked_pts_tensor(scene_pts, pix, vis, masks[off[i]:off[i + 1]].astype(np.float32))
        np.testing.assert_array_equal(rows, rows_g[off[i]:off[i + 1]])


def test_mask_functions(cuda_dev, G):
    P = _load(os.path.join(ROOT, "tools", "projection_2d_to_3d.py"), "bff_tools_projection")
    R = _load(os.path.join(ROOT, "tools", "refinement.py"), "bff_tools_refinement")
    a, b = torch.from_numpy(_bits(G["iou_a"], 1501)), torch.from_numpy(_bits(G["iou_b"], 1501))
    iou = P.calculate_iou(a).numpy()
    nan = np.isnan(G["iou_aa"])
    assert np.array_equal(np.isnan(iou), nan)
    np.testing.assert_array_equal(iou[~nan].view(np.int32), G["iou_aa"][~nan].view(np.int32))
    got = R.calculate_iou_between_stages(a.to(torch.uint8), b)
    assert got.shape == (7, 23)
    np.testing.assert_array_equal(got.numpy().view(np.int32), G["iou_ab_T"].view(np.int32))
    comps = [[int(x) for x in s.split(",")] if s else [] for s in G["components"].tolist()]
    assert P.find_unconnected_subgraphs_tensor(torch.from_numpy(G["adj"]).float()) == comps
    groups = [[int(x) for x in s.split(",")] for s in G["overlap_groups"].tolist()]
    out = P.solve_overlapping(torch.from_numpy(_bits(G["overlap_in"], 1501)), groups)
    np.testing.assert_array_equal(out.numpy(), _bits(G["overlap_out"], 1501))
    hostile = {"length": 100, "counts": G["rle_hostile_counts"]}
    np.testing.assert_array_equal(R.rle_decode(hostile), G["rle_hostile_decoded"])
    # aggregate(): groups + merged rows + sequential-mean confidences, against the dense formulation
    conf = torch.linspace(0.2, 0.45, a.shape[0])
    res, lists = P.aggregate({"ins": a, "conf": conf, "final_class": ["x"] * a.shape[0]}, 0.12, 0.75)
    kept = [c for c in comps if len(c) >= 2]
    assert lists == kept and res["final_class"] == ["x"] * len(kept)
    np.testing.assert_array_equal(res["ins"].cpu().numpy(), np.stack([a[c].any(0).numpy() for c in kept]))
    np.testing.assert_allclose(res["conf"].numpy(), [sum(conf[k] for k in c) / len(c) for c in kept], rtol=1e-6)
