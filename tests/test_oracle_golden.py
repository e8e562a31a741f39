"""CPU: pin the oracle (C restatement oracle/bff_oracle.c + numpy/torch port oracle/reference_port.py)
against golden vectors produced by the UNMODIFIED reference (oracle/make_golden.py, reference commit
1d3aff0f), and check the product's host-side logic against the same vectors. No GPU needed."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN


@pytest.fixture(scope="module")
def G():
    return np.load(os.path.join(GOLDEN, "functions.npz"), allow_pickle=False)


def _bits(packed, n):
    return np.unpackbits(packed, axis=1)[:, :n].astype(bool)


def _groups(arr):
    return [[int(x) for x in s.split(",")] if s else [] for s in arr.tolist()]


def test_c_oracle_projection_visibility_gather(G):
    import oracle as O
    xyz, K, depth = G["in_xyz"], G["in_K"], G["in_depth"]
    f, h, w = depth.shape
    masks = _bits(G["in_masks"], h * w).reshape(-1, h, w)
    off = G["in_mask_off"]
    rows_g = _bits(G["rows"], xyz.shape[0])
    assert G["vis"].sum() > 200 and rows_g.sum() > 50
    for i in range(f):
        cam, pix = O.project(xyz, G["pinv"][i], K)
        # camera-space coordinates: the BLAS FMA chain, bit for bit (NaN == NaN)
        np.testing.assert_array_equal(cam.view(np.int64), np.ascontiguousarray(G["cam"][i]).view(np.int64))
        np.testing.assert_array_equal(pix, G["pix"][i])
        vis = O.visibility(cam, pix, depth[i], 0.08)
        np.testing.assert_array_equal(vis, G["vis"][i])
        rows = O.mask_gather(pix, vis, masks[off[i]:off[i + 1]])
        np.testing.assert_array_equal(rows, rows_g[off[i]:off[i + 1]])
    rows, mc, vc = O.lift_scene(xyz, G["pinv"], K, depth, 0.08, masks.astype(np.uint8), off)
    np.testing.assert_array_equal(rows, rows_g)
    np.testing.assert_array_equal(mc[0], rows_g.sum(0))
    np.testing.assert_array_equal(vc, G["vis"].sum(0))


def test_port_projection_visibility_gather(G):
    import reference_port as RP
    xyz, K, depth = G["in_xyz"], G["in_K"], G["in_depth"]
    f, h, w = depth.shape
    masks = _bits(G["in_masks"], h * w).reshape(-1, h, w)
    off = G["in_mask_off"]
    rows_g = _bits(G["rows"], xyz.shape[0])
    pcd4 = RP.homogeneous_cloud(xyz)
    for i in range(f):
        with np.errstate(all="ignore"):
            cam = RP.camera_points(pcd4, G["in_poses"][i])
            pix = RP.project(cam, K)
            vis = RP.visibility(cam, pix, depth[i], 0.08)
        np.testing.assert_array_equal(pix, G["pix"][i])
        np.testing.assert_array_equal(vis, G["vis"][i])
        np.testing.assert_array_equal(RP.gather_masks(pix, vis, masks[off[i]:off[i + 1]]), rows_g[off[i]:off[i + 1]])


def test_iou_oracles(G):
    import oracle as O
    import reference_port as RP
    a, b = _bits(G["iou_a"], 1501), _bits(G["iou_b"], 1501)
    for got in (O.mask_iou(a, a)[1], RP.pairwise_iou(torch.from_numpy(a)).numpy()):
        nan = np.isnan(G["iou_aa"])
        assert nan.sum() == 4 and np.array_equal(np.isnan(got), nan)
        np.testing.assert_array_equal(got[~nan].view(np.int32), G["iou_aa"][~nan].view(np.int32))
    np.testing.assert_array_equal(O.mask_iou(a, b)[1].T.copy().view(np.int32), G["iou_ab_T"].view(np.int32))
    np.testing.assert_array_equal(RP.iou_between(torch.from_numpy(a).to(torch.uint8), torch.from_numpy(b)).numpy(), G["iou_ab_T"])


def test_components_and_overlaps(G):
    import reference_port as RP
    from bff_b200 import aggregate
    comps = _groups(G["components"])
    assert [] in comps and max(len(c) for c in comps) > 2
    adj = torch.from_numpy(G["adj"])
    assert RP.components(adj.float()) == comps
    assert aggregate.connected_groups(adj) == comps           # product host logic
    groups = _groups(G["overlap_groups"])
    ins = torch.from_numpy(_bits(G["overlap_in"], 1501))
    np.testing.assert_array_equal(RP.resolve_overlaps(ins.clone(), groups).numpy(), _bits(G["overlap_out"], 1501))
    # product: the pair program replayed densely gives the same rows
    inter = (ins.float() @ ins.float().T).to(torch.int32)
    exp = ins.clone()
    for dst, src in aggregate.overlap_program(inter, [len(g) for g in groups]):
        exp[dst] &= ~exp[src]
    np.testing.assert_array_equal(exp.numpy(), _bits(G["overlap_out"], 1501))


def test_rle_codec(G):
    import oracle as O
    import reference_port as RP
    from bff_b200 import rle
    m = G["rle_masks"]
    enc = rle.rle_encode_rows(torch.from_numpy(m))
    for i, s in enumerate(G["rle_counts"].tolist()):
        counts = np.array([int(x) for x in s.split(",")], np.int64)
        np.testing.assert_array_equal(enc[i]["counts"], counts)          # product encoder == reference encoder
        r = {"length": m.shape[1], "counts": counts}
        np.testing.assert_array_equal(O.rle_decode(r).astype(bool), m[i])
        np.testing.assert_array_equal(RP.rle_to_mask(r).astype(bool), m[i])
        lo, hi, _ = rle.runs_of(r)                                       # product decoder (host half)
        d = np.zeros(m.shape[1], bool)
        for a, b in zip(lo, hi):
            d[a:b] = True
        np.testing.assert_array_equal(d, m[i])
    hostile = {"length": 100, "counts": G["rle_hostile_counts"]}
    exp = G["rle_hostile_decoded"].astype(bool)
    np.testing.assert_array_equal(O.rle_decode(hostile).astype(bool), exp)
    np.testing.assert_array_equal(RP.rle_to_mask(hostile).astype(bool), exp)
    lo, hi, _ = rle.runs_of(hostile)
    d = np.zeros(100, bool)
    for a, b in zip(lo, hi):
        d[a:b] = True
    np.testing.assert_array_equal(d, exp)


# ---- depth ingest: cv2.resize(depth, (W, H)) (projection_2d_to_3d.py:436) restated
def test_resize_matches_golden():
    import oracle as O
    R = np.load(os.path.join(GOLDEN, "resize.npz"), allow_pickle=False)
    k = 0
    while f"raw{k}" in R.files:
        dst = R[f"dst{k}"]
        got = O.resize_bilinear(R[f"raw{k}"].astype(np.float32) / 1000, dst.shape[0], dst.shape[1])
        np.testing.assert_array_equal(got.view(np.int32), dst.view(np.int32))
        k += 1
    assert k == 5


@pytest.mark.parametrize("src_hw,dst_hw", [((480, 640), (968, 1296)), ((480, 640), (480, 640)), ((480, 640), (240, 320)),
                                           ((480, 640), (600, 800)), ((968, 1296), (480, 640)), ((480, 640), (481, 641)),
                                           ((480, 640), (1000, 1333)), ((120, 160), (200, 300)), ((30, 40), (60, 80)),
                                           ((480, 640), (96, 128)), ((7, 5), (19, 23))])
def test_resize_matches_cv2(src_hw, dst_hw):
    """The restatement against the reference's own call, live, bit for bit (cv2 of this machine)."""
    cv2 = pytest.importorskip("cv2")
    import oracle as O
    rng = np.random.default_rng(src_hw[0] * 7 + dst_hw[1])
    raw = rng.integers(0, 6000, src_hw).astype(np.uint16)
    raw[rng.random(src_hw) < 0.02] = 0
    depth = raw.astype(np.float32) / 1000
    want = cv2.resize(depth, (dst_hw[1], dst_hw[0]))
    got = O.resize_bilinear(depth, dst_hw[0], dst_hw[1])
    np.testing.assert_array_equal(got.view(np.int32), want.view(np.int32))
