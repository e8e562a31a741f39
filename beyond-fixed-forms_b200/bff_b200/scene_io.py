"""Readers/writers of the reference's on-disk tree. Same files, same
parsing calls as the reference so that the bytes handed to the kernels are the bytes the
reference computes on; everything after parsing goes to the GPU."""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

DEPTH_SCALE = 1000.0      # tools/projection_2d_to_3d.py:346 (hard-coded there, not a config key)
DEPTH_THRESH = 0.08       # tools/projection_2d_to_3d.py:438,565


def torch_load(path: str):
    """torch.load as the reference calls it (projection_2d_to_3d.py:396, refinement.py:182-183);
    the files hold numpy RLE arrays, so weights_only must be off on torch >= 2.6."""
    return torch.load(path, map_location="cpu", weights_only=False)


def load_cloud(scene_npy_dir: str, scene_id: str) -> np.ndarray:
    """(N,3) float64 xyz (projection_2d_to_3d.py:385-387; file written by tools/utils/ply2npy.py)."""
    return np.ascontiguousarray(np.load(os.path.join(scene_npy_dir, f"{scene_id}.npy"))[:, :3], dtype=np.float64)


def load_intrinsics(scene_2d_dir: str, scene_id: str) -> np.ndarray:
    """projection_2d_to_3d.py:370-376."""
    return np.loadtxt(os.path.join(scene_2d_dir, scene_id, "intrinsic", "intrinsic_color.txt"))[:3, :3]


def read_matrix4(path: str) -> np.ndarray:
    """np.loadtxt(path) for the 4 x 4 whitespace-separated text matrices of ScanNet pose files, ~9x faster (35 us instead
    of 300 us per file, and np.loadtxt holds the GIL, so the ingest thread pool did not hide it: 300 frames = 90 ms per
    scene). Python's float() and np.loadtxt both round decimal text correctly, so the doubles are the same bits (incl.
    inf / nan tokens); anything that is not exactly four lines of four plain tokens goes to np.loadtxt itself."""
    try:
        with open(path, "rb") as fh:
            lines = [ln.split() for ln in fh.read().splitlines()]
        lines = [ln for ln in lines if ln]
        if len(lines) == 4 and all(len(ln) == 4 for ln in lines):
            return np.array([[float(t) for t in ln] for ln in lines], dtype=np.float64)
    except ValueError:
        pass
    return np.loadtxt(path)


def load_pose_inverse(scene_2d_dir: str, scene_id: str, frame_id: str) -> np.ndarray:
    """np.linalg.inv(np.loadtxt(pose)) exactly as projection_2d_to_3d.py:422,425 (LAPACK on the
    host; a rigid-transform shortcut would give different bits). Non-finite poses give NaNs."""
    pose = read_matrix4(os.path.join(scene_2d_dir, scene_id, "pose", f"{frame_id}.txt"))
    return np.linalg.inv(pose)


def list_view_frames(scene_2d_dir: str, scene_id: str, downsample_ratio: int) -> List[str]:
    """Frame ids of color/*.jpg sorted numerically, every `downsample_ratio`-th
    (projection_2d_to_3d.py:526-535)."""
    image_dir = os.path.join(scene_2d_dir, scene_id, "color")
    names = [f for f in os.listdir(image_dir) if f.endswith(".jpg")]
    names.sort(key=lambda x: int(x.split(".")[0]))
    return [f[:-4] for f in names[::downsample_ratio]]


def load_depth_raw(scene_2d_dir: str, scene_id: str, frame_id: str) -> np.ndarray:
    import cv2
    path = os.path.join(scene_2d_dir, scene_id, "depth", f"{frame_id}.png")
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise FileNotFoundError(path)      # the reference dies on None.astype (AttributeError)
    return img


def depth_to_metres(raw: np.ndarray, width: int, height: int) -> np.ndarray:
    """float32 depth at mask resolution as projection_2d_to_3d.py:431-436 builds it (host path,
    used when a resize is needed; OpenCV bilinear kept on the host -- SURVEY.md 8f N2)."""
    import cv2
    d = raw.astype(np.float32) / int(DEPTH_SCALE)
    return cv2.resize(d, (int(width), int(height)))


class FrameSet:
    """Frames of one scene ready for the device: inverse poses, depth stack, viewed flags.
    native_size = (Hs, Ws) when `depth` is raw uint16 at the sensor's size and the lift kernel resamples it
    to mask resolution itself (BFF_DEPTH_U16_BILINEAR); None when depth already has mask resolution."""

    def __init__(self, frame_ids: List[str], pose_inv: np.ndarray, depth: np.ndarray, depth_is_raw: bool,
                 viewed: np.ndarray, depth_range: Optional[Tuple[float, float]], native_size=None):
        self.frame_ids, self.pose_inv, self.depth = frame_ids, pose_inv, depth
        self.depth_is_raw, self.viewed, self.depth_range = depth_is_raw, viewed, depth_range
        self.native_size = native_size


INGEST_THREADS = int(os.environ.get("BFF_INGEST_THREADS", str(min(32, os.cpu_count() or 1))))   # frame files read in parallel
IN_KERNEL_RESIZE = True      # False: always resize on the host with cv2.resize like the reference (:436)
_resize_checked = {}


def device_resize_matches_cv2(src_hw, dst_hw, device) -> bool:
    """Run-time gate of the in-kernel depth resize: bff_depth_resize (the arithmetic the lift kernel uses for
    BFF_DEPTH_U16_BILINEAR) against THIS machine's cv2.resize on a seeded probe image of the same size pair,
    bit for bit. cv2's result depends on its build (IPP or not); on a mismatch the frames are resized on the
    host by cv2 itself, as the reference does, and handed over as float32. Checked once per size pair."""
    key = (tuple(src_hw), tuple(dst_hw), str(device))
    if key not in _resize_checked:
        import cv2
        from . import ops
        rng = np.random.default_rng(12345)
        raw = rng.integers(300, 9000, tuple(src_hw)).astype(np.uint16)
        raw[rng.random(tuple(src_hw)) < 0.03] = 0
        want = depth_to_metres(raw, dst_hw[1], dst_hw[0])
        got = ops.depth_resize(torch.from_numpy(raw.view(np.int16)[None]).to(device), dst_hw[0], dst_hw[1], DEPTH_SCALE)[0].cpu().numpy()
        _resize_checked[key] = bool(np.array_equal(got.view(np.int32), want.view(np.int32)))
    return _resize_checked[key]


def load_frames(scene_2d_dir: str, scene_id: str, frame_ids: Sequence[str], viewed: Sequence[bool],
                width: int, height: int, device=None) -> FrameSet:
    f = len(frame_ids)
    pose_inv = np.empty((f, 4, 4), np.float64)

    # PNG decoding (cv2 releases the GIL) dominates the wall time of a scene once the kernels take milliseconds: the
    # frames are decoded on a thread pool, in the order of frame_ids. The pose files are parsed on THIS thread meanwhile:
    # text parsing holds the GIL, and inside the pool 300 poses took 66-220 ms of contended time instead of 12 ms.
    workers = min(INGEST_THREADS, f)
    if workers > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=workers) as ex:
            pending = [ex.submit(load_depth_raw, scene_2d_dir, scene_id, fid) for fid in frame_ids]
            for i, fid in enumerate(frame_ids):
                pose_inv[i] = load_pose_inverse(scene_2d_dir, scene_id, fid)
            raws = [p.result() for p in pending]
    else:
        raws = []
        for i, fid in enumerate(frame_ids):
            pose_inv[i] = load_pose_inverse(scene_2d_dir, scene_id, fid)
            raws.append(load_depth_raw(scene_2d_dir, scene_id, fid))
    same = all(r.dtype == np.uint16 and r.shape == (height, width) for r in raws)
    if same:
        depth = np.stack(raws) if raws else np.zeros((0, height, width), np.uint16)
        rng = (0.0, 65535.0 / DEPTH_SCALE)
        return FrameSet(list(frame_ids), pose_inv, depth, True, np.asarray(viewed, np.uint8), rng)
    native = raws[0].shape if raws else None
    if (IN_KERNEL_RESIZE and device is not None and raws and all(r.dtype == np.uint16 and r.ndim == 2 and r.shape == native for r in raws)
            and device_resize_matches_cv2(native, (height, width), device)):
        # blended values lie between 0 and the largest sample: the culling window starts at 0
        return FrameSet(list(frame_ids), pose_inv, np.stack(raws), True, np.asarray(viewed, np.uint8),
                        (0.0, 65535.0 / DEPTH_SCALE), native_size=(int(native[0]), int(native[1])))
    depth = np.stack([depth_to_metres(r, width, height) for r in raws]) if raws else np.zeros((0, height, width), np.float32)
    rng = None
    if depth.size and np.isfinite(depth).all():
        rng = (float(min(depth.min(), 0.0)), float(depth.max()))
    return FrameSet(list(frame_ids), pose_inv, depth, False, np.asarray(viewed, np.uint8), rng)


def cloud_for_device(xyz: np.ndarray, device) -> torch.Tensor:
    """float4 fp32 when the float64 cloud is float32-representable (ScanNet PLYs are), else (N,3) f64."""
    x32 = xyz.astype(np.float32)
    if np.array_equal(x32.astype(np.float64), xyz):
        x4 = np.zeros((xyz.shape[0], 4), np.float32)
        x4[:, :3] = x32
        return torch.from_numpy(x4).to(device)
    return torch.from_numpy(np.ascontiguousarray(xyz, np.float64)).to(device)


def empty_stage2() -> dict:
    """What the reference saves when nothing is lifted/aggregated (projection_2d_to_3d.py:465-478,496-509)."""
    return {"ins": torch.tensor([[]]), "conf": torch.tensor([]), "final_class": []}


def save_result(result: dict, out_dir: str, cls: str, scene_id: str):
    os.makedirs(os.path.join(out_dir, cls), exist_ok=True)
    torch.save(result, os.path.join(out_dir, cls, f"{scene_id}.pth"))


def read_checkpoint(path: str) -> dict:
    import yaml
    if os.path.exists(path):
        with open(path, "r") as fh:
            return yaml.safe_load(fh) or {}
    return {}


def write_checkpoint(path: str, ckpt: dict):
    import yaml
    with open(path, "w") as fh:
        yaml.safe_dump(ckpt, fh)
