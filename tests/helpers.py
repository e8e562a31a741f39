"""Shared builders of seeded test inputs (small scenes in memory)."""
import numpy as np


def look_at(eye, target):
    fwd = target - eye
    fwd = fwd / np.linalg.norm(fwd)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return c2w


def pack_rows(dense):
    """(R,N) bool -> (R, ceil(N/32)) uint32, bit b of word w = element 32*w+b (numpy reference)."""
    dense = np.asarray(dense, dtype=bool)
    r, n = dense.shape
    wd = (n + 31) // 32
    pad = np.zeros((r, wd * 32), dtype=bool)
    pad[:, :n] = dense
    bits = np.packbits(pad.reshape(r, wd, 32), axis=2, bitorder="little")
    return bits.reshape(r, wd, 4).copy().view("<u4").reshape(r, wd)


def unpack_rows(packed, n):
    packed = np.ascontiguousarray(packed).view("<u4")
    r = packed.shape[0]
    bits = np.unpackbits(packed.view(np.uint8).reshape(r, -1), axis=1, bitorder="little")
    return bits[:, :n].astype(bool)


def random_scene(seed, n=5000, f=7, h=60, w=80, m_per_frame=(0, 1, 2, 3), n_queries=1, f32_xyz=True):
    """A box room seen by a few cameras with a depth image that matches a good part of the
    points (so that visibility is non-trivial), random blob masks, plus hostile values."""
    rng = np.random.default_rng(seed)
    xyz = rng.uniform([0, 0, 0], [6, 5, 2.8], size=(n, 3))
    if f32_xyz:
        xyz = xyz.astype(np.float32).astype(np.float64)
    fx = 577.870605 * w / 640.0
    K = np.array([[fx, 0.0, (w - 1) / 2.0], [0.0, fx, (h - 1) / 2.0], [0.0, 0.0, 1.0]])
    poses = np.empty((f, 4, 4))
    depth = np.zeros((f, h, w), np.float32)
    for i in range(f):
        eye = rng.uniform([1, 1, 1], [5, 4, 1.8])
        tgt = rng.uniform([0, 0, 0], [6, 5, 2])
        poses[i] = look_at(eye, tgt)
        pinv = np.linalg.inv(poses[i])
        cam = (pinv @ np.concatenate([xyz, np.ones((n, 1))], 1).T).T[:, :3]
        with np.errstate(all="ignore"):
            uv = (K @ cam.T / cam[:, 2]).T[:, :2]
        px = np.round(uv).astype(np.int64)
        ok = (cam[:, 2] > 0.1) & (px[:, 0] >= 0) & (px[:, 0] < w) & (px[:, 1] >= 0) & (px[:, 1] < h)
        # z-buffer of a random half of the points, quantised to millimetres like a sensor
        sel = np.nonzero(ok & (rng.random(n) < 0.5))[0]
        order = sel[np.argsort(-cam[sel, 2])]
        dimg = np.zeros((h, w))
        dimg[px[order, 1], px[order, 0]] = cam[order, 2]
        dimg = np.round(dimg * 1000.0)
        dimg += rng.integers(-90, 91, size=dimg.shape) * (dimg > 0)     # straddle the 0.08 m threshold
        dimg[rng.random((h, w)) < 0.05] = 0
        depth[i] = (np.clip(dimg, 0, 65535).astype(np.uint16).astype(np.float32) / 1000)
    masks, off, query = [], [0], []
    yy, xx = np.mgrid[0:h, 0:w]
    for i in range(f):
        m = int(rng.choice(m_per_frame))
        for _ in range(m):
            cy, cx, r = rng.uniform(0, h), rng.uniform(0, w), rng.uniform(3, max(h, w) / 2.5)
            masks.append(((yy - cy) ** 2 + (xx - cx) ** 2) < r * r)
            query.append(int(rng.integers(0, n_queries)))
        off.append(len(masks))
    masks = np.stack(masks) if masks else np.zeros((0, h, w), bool)
    return dict(xyz=xyz, K=K, poses=poses, pinv=np.stack([np.linalg.inv(p) for p in poses]), depth=depth,
                masks=masks, mask_off=np.asarray(off, np.int32), mask_query=np.asarray(query, np.int32),
                h=h, w=w, n=n, f=f, n_queries=n_queries)


# ----------------------------------------------------------------------------- golden tree helpers
def load_expected(name="e2e.pth"):
    import os
    import torch
    from conftest import GOLDEN
    return torch.load(os.path.join(GOLDEN, name), weights_only=False)


def copy_tree(tmp_path, name="tree"):
    """A writable copy of tests/golden/<name> (the CLIs write outputs and checkpoints under cwd)."""
    import os
    import shutil
    from conftest import GOLDEN
    dst = os.path.join(str(tmp_path), name)
    shutil.copytree(os.path.join(GOLDEN, name), dst)
    os.makedirs(os.path.join(dst, "checkpoints"), exist_ok=True)
    return dst


def fake_clip_sim(query, label):
    """What refinement.py:93-117 computes with the fake `clip` of oracle/run_reference.py."""
    import run_reference as RR
    a, b = RR.fake_clip_similarity_vector(query), RR.fake_clip_similarity_vector(label)
    sim = a @ b.T
    sim = sim / (a.norm(dim=-1, keepdim=True) * b.norm(dim=-1, keepdim=True).T)
    return sim.item()


def check_against_expected(result, rec, what):
    """result: dict as saved; rec: record of oracle/make_golden.py:_pack_result."""
    import torch
    assert list(result["final_class"]) == rec["final_class"], what
    if rec["ins_dtype"] == "list":
        assert isinstance(result["ins"], list) and len(result["ins"]) == 0, what
        assert isinstance(result["conf"], list) and len(result["conf"]) == 0, what
        return
    ins = result["ins"]
    assert isinstance(ins, torch.Tensor) and tuple(ins.shape) == rec["ins_shape"], (what, ins.shape, rec["ins_shape"])
    assert str(ins.dtype) == rec["ins_dtype"], (what, ins.dtype)
    if ins.numel():
        exp = np.unpackbits(rec["ins_bits"], axis=1)[:, :ins.shape[1]].astype(bool)
        np.testing.assert_array_equal(ins.cpu().numpy().astype(bool), exp, err_msg=what)       # bit-exact masks
    conf = result["conf"]
    assert isinstance(conf, torch.Tensor) == rec["conf_is_tensor"], what
    np.testing.assert_allclose(np.asarray(conf, dtype=np.float32), np.asarray(rec["conf"], dtype=np.float32),
                               rtol=1e-5, err_msg=what)                                      # north_star tolerance


def port_run_class(cfg, cls):
    """oracle/reference_port.py driven over an on-disk tree like the reference's two CLIs.
    Returns ({scene: stage-2 dict}, {scene: final dict}). cwd must be the tree root."""
    import os
    import torch
    import reference_port as RP
    from bff_b200 import scene_io
    from bff_b200.refine import STAGE1_CLASS_NAMES
    h, w = cfg["height_2d"], cfg["width_2d"]
    seg_dir = os.path.join(cfg["mask_2d_dir"], cls)
    stage2, finals, p1s = {}, {}, []
    scenes = sorted(s[:-4] for s in os.listdir(seg_dir) if s.endswith("_00.pth"))
    for sid in scenes:
        K = scene_io.load_intrinsics(cfg["scene_2d_dir"], sid)
        xyz = scene_io.load_cloud(cfg["scene_npy_dir"], sid)

        def frame(fid):
            pose = np.loadtxt(os.path.join(cfg["scene_2d_dir"], sid, "pose", f"{fid}.txt"))
            depth = scene_io.depth_to_metres(scene_io.load_depth_raw(cfg["scene_2d_dir"], sid, fid), w, h)
            return {"pose": pose, "depth": depth}
        mask_frames = []
        for r in scene_io.torch_load(os.path.join(seg_dir, f"{sid}.pth")):
            fr = frame(r["frame_id"][:-4])
            fr["masks"] = np.stack([RP.rle_to_mask(m).reshape(h, w) for m in r["segmented_frame_masks"]]).astype(np.float32)
            fr["conf"], fr["labels"] = r["confidences"], r["labels"]
            mask_frames.append(fr)
        view = [frame(f) for f in scene_io.list_view_frames(cfg["scene_2d_dir"], sid, cfg["downsample_ratio"])]
        stage2[sid] = RP.stage2_scene(xyz, K, mask_frames, view, cfg)
        s1 = scene_io.torch_load(os.path.join(cfg["stage_1_results_dir"], f"{sid}.pth"))
        s1_ins = torch.from_numpy(np.stack([RP.rle_to_mask(r) for r in s1["ins"]]))
        labels = [STAGE1_CLASS_NAMES[int(i)] for i in s1["final_class"]]
        p1s.append(RP.refine_pass1(s1_ins, labels, stage2[sid], cls, cfg["stage1_iou_thres"], fake_clip_sim))
    thr = RP.sim_threshold([p["sims"] for p in p1s], cfg["refinment_sim_percentile"])
    for sid, p1 in zip(scenes, p1s):
        finals[sid] = RP.refine_pass2(p1, cls, cfg["refiment_iou_thres"], thr)
    return stage2, finals
