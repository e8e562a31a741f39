"""Seeded synthetic ScanNet200-shaped scenes.

Inputs of the hot path are produced by models and sensors that are out of scope (2D
segmentation, RGB-D capture); this module fabricates them with the shapes and on-disk
formats the reference reads (tools/projection_2d_to_3d.py:370-400,422,431-436,526-535;
tools/refinement.py:172-193): a box room with box "objects", points sampled on the
surfaces (float32-representable, stored float64 like tools/utils/ply2npy.py:9-20),
a looped camera trajectory, analytic ray/AABB depth in millimetres (uint16) and
object-consistent 2D instance masks per text query.

Ray casting uses torch tensor ops so it can run on the GPU for the big bench scenes; it
is input fabrication, not part of the measured path.
"""
from __future__ import annotations

import dataclasses
import math
import os
from typing import Iterator, List, Optional

import numpy as np
import torch

SCANNET200_STAGE1_NAMES = None  # filled lazily from .refine (198 names, refinement.py:58-62)


@dataclasses.dataclass
class SceneSpec:
    n_points: int = 20000
    n_frames: int = 12            # frames listed in color/ AFTER down-sampling (F_all)
    width: int = 320              # mask / colour resolution (cfg.width_2d)
    height: int = 240
    depth_width: int = 0          # native depth resolution (0 -> same as mask resolution)
    depth_height: int = 0
    n_queries: int = 1
    n_boxes: int = 12
    room: tuple = (6.0, 5.0, 2.8)
    n_stage1: int = 20
    downsample_ratio: int = 10
    mask_drop_prob: float = 0.3
    seed: int = 0


@dataclasses.dataclass
class Scene:
    spec: SceneSpec
    xyz: np.ndarray               # (N,3) float64, float32-representable
    rgb: np.ndarray               # (N,3) float64 in [0,255]
    point_obj: np.ndarray         # (N,) int32, -1 = room shell, b = box id
    K: np.ndarray                 # (3,3) float64 at mask resolution
    poses: np.ndarray             # (F,4,4) float64 camera->world
    frame_ids: List[int]          # file stems of the F_all frames (multiples of downsample_ratio)
    box_lo: np.ndarray            # (B,3)
    box_hi: np.ndarray            # (B,3)
    box_query: np.ndarray         # (B,) int32 query id or -1


def _look_at(eye, target):
    """camera->world with x right, y down, z forward (ScanNet/OpenCV convention)."""
    fwd = target - eye
    fwd = fwd / np.linalg.norm(fwd)
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return c2w


def make_scene(spec: SceneSpec) -> Scene:
    rng = np.random.default_rng(spec.seed)
    Lx, Ly, Lz = spec.room
    B = spec.n_boxes
    size = rng.uniform([0.3, 0.3, 0.3], [1.2, 1.2, 1.4], size=(B, 3))
    lo = np.empty((B, 3))
    lo[:, 0] = rng.uniform(0.2, Lx - 0.2 - size[:, 0])
    lo[:, 1] = rng.uniform(0.2, Ly - 0.2 - size[:, 1])
    lo[:, 2] = 0.0
    hi = lo + size
    hi[:, 2] = np.minimum(hi[:, 2], Lz - 0.2)
    box_query = (np.arange(B) % (spec.n_queries + 1)).astype(np.int32) - 1  # -1: distractor
    if spec.n_queries == 1:
        box_query = np.where(np.arange(B) % 3 == 2, -1, 0).astype(np.int32)

    # --- surface sampling, area weighted: 6 room faces + 5 faces per box (no bottom)
    faces = []  # (obj, axis, const, (u_lo,u_hi), (v_lo,v_hi))
    room_lo, room_hi = np.zeros(3), np.array([Lx, Ly, Lz])
    for ax in range(3):
        for c in (room_lo[ax], room_hi[ax]):
            faces.append((-1, ax, c, room_lo, room_hi))
    for b in range(B):
        for ax in range(3):
            for c in (lo[b, ax], hi[b, ax]):
                if ax == 2 and c == lo[b, 2]:
                    continue
                faces.append((b, ax, c, lo[b], hi[b]))
    areas = []
    for (_o, ax, _c, l, h) in faces:
        o = [a for a in range(3) if a != ax]
        areas.append((h[o[0]] - l[o[0]]) * (h[o[1]] - l[o[1]]))
    areas = np.asarray(areas)
    which = rng.choice(len(faces), size=spec.n_points, p=areas / areas.sum())
    which.sort()  # face-major order: a mesh-like, spatially coherent vertex order
    xyz = np.empty((spec.n_points, 3))
    pobj = np.empty(spec.n_points, np.int32)
    for fi, (o_id, ax, c, l, h) in enumerate(faces):
        sel = np.nonzero(which == fi)[0]
        if sel.size == 0:
            continue
        p = rng.uniform(l, h, size=(sel.size, 3))
        p[:, ax] = c
        # scan-line order inside a face keeps neighbours close like mesh vertices
        o = [a for a in range(3) if a != ax]
        cell = np.floor(p[:, o[0]] / 0.25) * 1000 + np.floor(p[:, o[1]] / 0.25)
        p = p[np.argsort(cell, kind="stable")]
        xyz[sel] = p
        pobj[sel] = o_id
    xyz += rng.normal(0, 0.004, xyz.shape)          # sensor-ish noise
    xyz = xyz.astype(np.float32).astype(np.float64)  # float32-representable (PLY is float32)
    rgb = np.floor(rng.uniform(0, 256, size=(spec.n_points, 3)))

    W, H = spec.width, spec.height
    fx = 577.870605 * W / 640.0
    K = np.array([[fx, 0.0, (W - 1) / 2.0], [0.0, fx, (H - 1) / 2.0], [0.0, 0.0, 1.0]])

    F = spec.n_frames
    poses = np.empty((F, 4, 4))
    ctr = np.array([Lx / 2, Ly / 2, 1.4])
    for f in range(F):
        a = 2 * math.pi * f / max(F, 1) + rng.normal(0, 0.02)
        eye = ctr + np.array([0.28 * Lx * math.cos(a), 0.28 * Ly * math.sin(a), rng.uniform(-0.2, 0.2)])
        tgt = ctr + np.array([0.45 * Lx * math.cos(a + 2.4), 0.45 * Ly * math.sin(a + 2.4), rng.uniform(-0.9, -0.2)])
        poses[f] = _look_at(eye, tgt)
    frame_ids = [f * spec.downsample_ratio for f in range(F)]
    return Scene(spec, xyz, rgb, pobj, K, poses, frame_ids, lo, hi, box_query)


def _raycast(scene: Scene, c2w: np.ndarray, K: np.ndarray, W: int, H: int, device):
    """Analytic ray/AABB cast. Returns (z_depth (H,W) float32 metres, obj (H,W) int16)."""
    dt = torch.float32
    Kt = torch.as_tensor(K, dtype=dt, device=device)
    u = torch.arange(W, device=device, dtype=dt)
    v = torch.arange(H, device=device, dtype=dt)
    dx = (u - Kt[0, 2]) / Kt[0, 0]
    dy = (v - Kt[1, 2]) / Kt[1, 1]
    dcam = torch.stack([dx[None, :].expand(H, W), dy[:, None].expand(H, W), torch.ones(H, W, device=device, dtype=dt)], -1)
    R = torch.as_tensor(c2w[:3, :3], dtype=dt, device=device)
    o = torch.as_tensor(c2w[:3, 3], dtype=dt, device=device)
    d = dcam.reshape(-1, 3) @ R.T                                   # (P,3) world dirs, camera z == 1
    d = torch.where(d.abs() < 1e-9, torch.full_like(d, 1e-9), d)
    inv = 1.0 / d
    lo = torch.as_tensor(scene.box_lo, dtype=dt, device=device)     # (B,3)
    hi = torch.as_tensor(scene.box_hi, dtype=dt, device=device)
    t0 = (lo[None] - o) * inv[:, None, :]
    t1 = (hi[None] - o) * inv[:, None, :]
    tn = torch.minimum(t0, t1).amax(-1)                             # (P,B)
    tf = torch.maximum(t0, t1).amin(-1)
    hit = (tf >= tn) & (tn > 1e-3)
    tbox = torch.where(hit, tn, torch.full_like(tn, float("inf")))
    tb, ib = tbox.min(dim=1)
    room_hi = torch.as_tensor(scene.spec.room, dtype=dt, device=device)
    r0 = (0.0 - o) * inv
    r1 = (room_hi - o) * inv
    troom = torch.maximum(r0, r1).amin(-1)                          # exit distance (camera is inside)
    t = torch.minimum(tb, troom)
    obj = torch.where(tb < troom, ib, torch.full_like(ib, -1)).to(torch.int16)
    return t.reshape(H, W), obj.reshape(H, W)


def render_frame(scene: Scene, f: int, device="cpu", gen: Optional[torch.Generator] = None):
    """Depth in millimetres (uint16 as int32 tensor, native depth resolution) and the object-id
    buffer at MASK resolution for frame f."""
    sp = scene.spec
    W, H = sp.width, sp.height
    Wd, Hd = (sp.depth_width or W), (sp.depth_height or H)
    if (Wd, Hd) == (W, H):
        z, obj = _raycast(scene, scene.poses[f], scene.K, W, H, device)
        zd = z
    else:
        _, obj = _raycast(scene, scene.poses[f], scene.K, W, H, device)
        Kd = scene.K.copy()
        Kd[0] *= Wd / W
        Kd[1] *= Hd / H
        zd, _ = _raycast(scene, scene.poses[f], Kd, Wd, Hd, device)
    if gen is None:
        gen = torch.Generator(device=device).manual_seed(sp.seed * 100003 + f)
    noise = (torch.rand(zd.shape, generator=gen, device=device) - 0.5) * 6.0      # +-3 mm
    mm = torch.clamp(torch.round(zd * 1000.0 + noise), 0, 65535)
    holes = torch.rand(zd.shape, generator=gen, device=device) < 0.02
    mm = torch.where(holes | ~torch.isfinite(zd), torch.zeros_like(mm), mm).to(torch.int32)
    return mm, obj


def frame_masks(scene: Scene, f: int, obj: torch.Tensor, query: int, rng: np.random.Generator, min_pixels: int = 30):
    """2D instance masks (list of (H,W) bool tensors) of one query in one frame: one per visible
    box of that query, randomly dropped and jittered by up to 2 px (imperfect segmentation)."""
    out = []
    for b in np.nonzero(scene.box_query == query)[0]:
        m = obj == int(b)
        if int(m.sum()) < min_pixels or rng.random() < scene.spec.mask_drop_prob:
            continue
        sx, sy = (int(s) for s in rng.integers(-2, 3, size=2))
        out.append(torch.roll(m, shifts=(sy, sx), dims=(0, 1)))
    return out


def stage1_masks(scene: Scene, query_names: List[str], class_names: List[str], rng: np.random.Generator):
    """Stage-1 (Open3DIS-like) proposals: (n,N) bool masks, class indices into the stage-1
    name table (refinement.py:58-62) and confidences."""
    sp = scene.spec
    N = sp.n_points
    masks, cls = [], []
    boxes = list(range(sp.n_boxes))
    k = 0
    while len(masks) < sp.n_stage1:
        if k < len(boxes):
            b = boxes[k]
            m = scene.point_obj == b
            flip = rng.random(N) < 0.1
            m = np.where(flip & m, False, m)
            grow = (rng.random(N) < 0.002)
            m = m | grow
            q = scene.box_query[b]
            name = query_names[q].replace(" ", "_") if q >= 0 else None
            ci = class_names.index(name) if name in class_names else int(rng.integers(0, len(class_names)))
        else:
            # background slab: a contiguous index range (superpoint-like)
            a = int(rng.integers(0, max(N - 10, 1)))
            ln = int(rng.integers(5, max(N // 20, 6)))
            m = np.zeros(N, bool)
            m[a:a + ln] = True
            ci = int(rng.integers(0, len(class_names)))
            if k % 7 == 0 and query_names:
                nm = query_names[0].replace(" ", "_")
                ci = class_names.index(nm) if nm in class_names else ci
        masks.append(m)
        cls.append(ci)
        k += 1
    return np.stack(masks), np.asarray(cls, np.int64), np.ones(len(masks), np.float32)


DEFAULT_CFG = {
    # keys consumed by the hot path (configs/config.yaml of the reference)
    "downsample_ratio": 10,
    "iou_thres": 0.2,
    "similarity_thres": 0.75,
    "min_aggragated_masks": 2,
    "if_occurance_threshold": False,
    "occurance_threshold": 0.3,
    "if_detected_ratio_threshold": True,
    "detected_ratio_threshold": 0.38,
    "remove_filtered_masks": 0.4,
    "remove_small_masks": 5,
    "stage1_iou_thres": 0.1,
    "refinment_sim_thres": 0.65,
    "refinment_sim_percentile": 0.2,
    "refiment_iou_thres": 0.45,
}


def write_tree(root: str, scenes: dict, queries: List[str], cfg_overrides: Optional[dict] = None, device="cpu") -> str:
    """Write the reference's on-disk input tree for `scenes`
    ({scene_id: Scene}) and return the path of the config yaml."""
    import cv2
    import yaml
    from .refine import STAGE1_CLASS_NAMES
    from .rle import rle_encode_rows

    p = lambda *a: os.path.join(root, *a)
    first = next(iter(scenes.values()))
    cfg = dict(DEFAULT_CFG)
    cfg.update({
        "scene_npy_dir": p("data", "Scannet200", "Scannet200_3D", "original_npy_files"),
        "scene_2d_dir": p("data", "Scannet200", "Scannet200_2D"),
        "mask_2d_dir": p("output", "mask_2d", "rle"),
        "mask_3d_dir": p("output", "mask_3d"),
        "stage_1_results_dir": p("exp_stage_1"),
        "final_output_dir": p("output", "final_output"),
        "width_2d": first.spec.width,
        "height_2d": first.spec.height,
        "downsample_ratio": first.spec.downsample_ratio,
    })
    cfg.update(cfg_overrides or {})
    for d in ("scene_npy_dir", "mask_2d_dir", "mask_3d_dir", "stage_1_results_dir", "final_output_dir"):
        os.makedirs(cfg[d], exist_ok=True)
    os.makedirs(p("checkpoints"), exist_ok=True)
    for scene_id, sc in scenes.items():
        sp = sc.spec
        rng = np.random.default_rng(sp.seed + 7)
        np.save(os.path.join(cfg["scene_npy_dir"], f"{scene_id}.npy"), np.concatenate([sc.xyz, sc.rgb], axis=1))
        s2d = os.path.join(cfg["scene_2d_dir"], scene_id)
        for sub in ("intrinsic", "pose", "depth", "color"):
            os.makedirs(os.path.join(s2d, sub), exist_ok=True)
        K4 = np.eye(4)
        K4[:3, :3] = sc.K
        np.savetxt(os.path.join(s2d, "intrinsic", "intrinsic_color.txt"), K4)
        for n in range(sc.frame_ids[-1] + 1):
            open(os.path.join(s2d, "color", f"{n}.jpg"), "wb").close()      # only NAMES are read
        per_query = {q: [] for q in queries}
        for f, fid in enumerate(sc.frame_ids):
            np.savetxt(os.path.join(s2d, "pose", f"{fid}.txt"), sc.poses[f])
            mm, obj = render_frame(sc, f, device)
            cv2.imwrite(os.path.join(s2d, "depth", f"{fid}.png"), mm.cpu().numpy().astype(np.uint16))
            for qi, q in enumerate(queries):
                ms = frame_masks(sc, f, obj.cpu(), qi, rng)
                if not ms:
                    continue                     # frames without boxes are absent (segmentation_2d.py:271-274)
                stack = torch.stack(ms).reshape(len(ms), -1)
                per_query[q].append({
                    "frame_id": f"{fid}.jpg",
                    "segmented_frame_masks": rle_encode_rows(stack),
                    "confidences": torch.as_tensor(rng.uniform(0.2, 0.45, len(ms)).astype(np.float32)),
                    "labels": [q] * len(ms),
                })
        for q in queries:
            os.makedirs(os.path.join(cfg["mask_2d_dir"], q), exist_ok=True)
            torch.save(per_query[q], os.path.join(cfg["mask_2d_dir"], q, f"{scene_id}.pth"))
        m1, c1, conf1 = stage1_masks(sc, queries, STAGE1_CLASS_NAMES, rng)
        torch.save({"ins": rle_encode_rows(torch.from_numpy(m1)), "conf": torch.from_numpy(conf1),
                    "final_class": torch.from_numpy(c1)},
                   os.path.join(cfg["stage_1_results_dir"], f"{scene_id}.pth"))
    cfg_path = p("config.yaml")
    with open(cfg_path, "w") as fh:
        yaml.safe_dump(cfg, fh)
    return cfg_path
