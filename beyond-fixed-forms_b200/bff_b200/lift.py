"""Scene lifting: host orchestration of bff_lift_scene (one launch per scene).

Mirrors the frame loops of tools/projection_2d_to_3d.py:413-461 (mask frames) and :538-567
(viewed-count frames), but visits each frame once: frames that carry masks and also count
into viewed_counts are flagged instead of being projected twice.
Supports Q concurrent text queries sharing projection and visibility."""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import ops, scene_io
from .rle import pairs_for_device


@dataclasses.dataclass
class QueryMasks:
    """Per-query contents of one <mask_2d_dir>/<cls>/<scene>.pth (projection_2d_to_3d.py:395-421)."""
    query: str
    frame_ids: List[str]                 # per listed frame, file order
    rles: List[list]                     # per frame: list of RLE dicts
    confidences: List[torch.Tensor]      # per frame: (M,) tensor
    labels: List[List[str]]              # per frame: list[str]

    @classmethod
    def from_file_records(cls, query: str, records: Sequence[dict]) -> "QueryMasks":
        return cls(query, [r["frame_id"][:-4] for r in records], [list(r["segmented_frame_masks"]) for r in records],
                   [r["confidences"] for r in records], [list(r["labels"]) for r in records])

    def flat_conf_and_labels(self):
        """Row-ordered confidences (list of 0-d tensors) and labels (projection_2d_to_3d.py:454-457)."""
        conf, lab = [], []
        for f in range(len(self.frame_ids)):
            for k in range(len(self.rles[f])):
                conf.append(self.confidences[f][k])
                lab.append(self.labels[f][k])
        return conf, lab


@dataclasses.dataclass
class LiftedScene:
    n_points: int
    rows: List[torch.Tensor]             # per query: (Ins_q, Wd) int32 packed rows, file order
    masked_counts: torch.Tensor          # (Q, N) int32
    viewed_counts: Optional[torch.Tensor]  # (N,) int32 or None
    n_frames: int


def plan_frames(queries: Sequence[QueryMasks], view_frames: Optional[Sequence[str]]):
    """Union of frames to visit. Returns (frame_ids, viewed flags, per-frame list of
    (query index, frame index inside that query))."""
    order: List[str] = []
    slot: Dict[str, int] = {}
    per_frame: List[list] = []
    for qi, q in enumerate(queries):
        seen_here = set()
        for fi, fid in enumerate(q.frame_ids):
            if fid in slot and fid not in seen_here:
                s = slot[fid]
            else:                          # new frame, or the same frame listed twice by one file
                s = len(order)
                order.append(fid)
                per_frame.append([])
                slot.setdefault(fid, s)
            seen_here.add(fid)
            per_frame[s].append((qi, fi))
    viewed = [False] * len(order)
    if view_frames is not None:
        done = set()
        for fid in view_frames:
            if fid in done:
                continue
            done.add(fid)
            if fid in slot:
                viewed[slot[fid]] = True
            else:
                slot[fid] = len(order)
                order.append(fid)
                per_frame.append([])
                viewed.append(True)
    return order, viewed, per_frame


def mask_layout(per_frame: List[list], queries: Sequence[QueryMasks], lo: int, hi: int):
    """Masks of frame slots [lo, hi) in launch order: (rles, query id per mask, per-frame offsets)."""
    rles, mask_query, offsets = [], [], [0]
    for s in range(lo, hi):
        for (qi, fi) in per_frame[s]:
            for rle in queries[qi].rles[fi]:
                rles.append(rle)
                mask_query.append(qi)
        offsets.append(len(rles))
    return rles, mask_query, offsets


def query_row_index(per_frame: List[list], queries: Sequence[QueryMasks]) -> List[List[int]]:
    """For each query, the positions of its rows in the launch layout, in the query's own file
    order (frame order of its .pth, mask order inside a frame: projection_2d_to_3d.py:454-457)."""
    pos = [dict() for _ in queries]
    g = 0
    for s in range(len(per_frame)):
        for (qi, fi) in per_frame[s]:
            for k in range(len(queries[qi].rles[fi])):
                pos[qi][(fi, k)] = g
                g += 1
    return [[pos[qi][(fi, k)] for fi in range(len(q.frame_ids)) for k in range(len(q.rles[fi]))]
            for qi, q in enumerate(queries)]


def lift_frames(xyz_dev: torch.Tensor, intrinsics: np.ndarray, frames: scene_io.FrameSet, rles: list,
                mask_query: List[int], offsets: List[int], n_queries: int, height: int, width: int,
                want_viewed: bool, depth_thresh: float = scene_io.DEPTH_THRESH, frames_per_chunk: int = 0) -> ops.LiftOutput:
    """RLE -> packed 2D masks on the device, then ONE bff_lift_scene launch over `frames`."""
    dev = xyz_dev.device
    hw = height * width
    masks2d = None
    if rles:
        counts, pair_off, length = pairs_for_device(rles, expect_length=None)
        if length != hw:
            raise ValueError(f"2D mask length {length} does not match height_2d*width_2d = {hw}")
        masks2d = ops.rle_pairs_to_bits_dev(torch.from_numpy(counts).to(dev), torch.from_numpy(pair_off).to(dev), hw)
    if frames.depth_is_raw:
        depth_dev = torch.from_numpy(frames.depth.view(np.int16)).to(dev)
    else:
        depth_dev = torch.from_numpy(np.ascontiguousarray(frames.depth, np.float32)).to(dev)
    return ops.lift_scene(
        xyz_dev, torch.from_numpy(frames.pose_inv).to(dev), intrinsics, depth_dev, height, width,
        masks2d=masks2d,
        mask_offset=torch.tensor(offsets, dtype=torch.int32, device=dev) if rles else None,
        mask_query=torch.tensor(mask_query, dtype=torch.int32, device=dev) if (rles and n_queries > 1) else None,
        n_queries=max(n_queries, 1),
        frame_viewed=torch.from_numpy(frames.viewed).to(dev) if want_viewed else None,
        depth_thresh=depth_thresh, depth_scale=scene_io.DEPTH_SCALE, depth_range=frames.depth_range,
        want_viewed=want_viewed, frames_per_chunk=frames_per_chunk, native_size=getattr(frames, "native_size", None))


def split_rows(all_rows: torch.Tensor, per_frame: List[list], queries: Sequence[QueryMasks]) -> List[torch.Tensor]:
    out = []
    for idx in query_row_index(per_frame, queries):
        if len(queries) == 1 and idx == list(range(all_rows.shape[0])):
            out.append(all_rows)
        else:
            out.append(all_rows[torch.tensor(idx, dtype=torch.long, device=all_rows.device)] if idx else all_rows[:0])
    return out


REORDER = "auto"            # "auto" | "always" | "never": lift on a Morton-sorted copy of the cloud (results are un-permuted)
REORDER_MIN_POINTS = 8192
REORDER_MIN_POINT_VIEWS = 2e8   # auto: check + sort + un-permute cost ~1.1 ms at 250k points / 525 rows (scripts/lift_point_order.py);
                                # an incoherent cloud costs the launch 2.8x, so sorting pays from ~2.5x the c2 scene upwards
REORDER_SPREAD = 0.2        # auto: mean extent of 32 consecutive points / extent of the cloud above which sorting pays


def point_order_spread(xyz_dev: torch.Tensor) -> float:
    """Mean bounding-box diagonal of groups of 32 consecutive points relative to the cloud's diagonal: ~0.04 for
    mesh-ordered ScanNet-like clouds, ~0.8 for shuffled ones. lift_kernel culls per WARP (32 consecutive points), so this
    number decides whether its frustum culling can work (measured at c2: 0.25 ms coherent, 0.70 ms shuffled)."""
    x = xyz_dev[:, :3]
    n32 = (x.shape[0] // 32) * 32
    if n32 == 0:
        return 0.0
    g = x[:n32].reshape(-1, 32, 3)
    ext = (g.max(dim=1).values - g.min(dim=1).values).double().norm(dim=1).mean()
    tot = (x.max(dim=0).values - x.min(dim=0).values).double().norm()
    r = float(ext / tot) if float(tot) > 0 else 0.0
    return r if r == r else 0.0            # NaN coordinates: leave the order alone


def morton_order(xyz_dev: torch.Tensor) -> torch.Tensor:
    """Permutation (int64, CUDA) that sorts the points along a 30-bit Morton curve of their bounding box."""
    x = xyz_dev[:, :3].double()
    lo, hi = x.min(dim=0).values, x.max(dim=0).values
    q = ((x - lo) / (hi - lo).clamp_min(1e-30) * 1023.0).clamp(0, 1023).long()

    def spread(v):
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        return (v | (v << 2)) & 0x09249249
    return torch.argsort(spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2), stable=True)


def lift(xyz_dev: torch.Tensor, intrinsics: np.ndarray, frames: scene_io.FrameSet, per_frame: List[list],
         queries: Sequence[QueryMasks], height: int, width: int, want_viewed: bool,
         depth_thresh: float = scene_io.DEPTH_THRESH, frames_per_chunk: int = 0) -> LiftedScene:
    """Single-GPU lifting of all planned frames; rows are returned per query in file order.
    When consecutive points of the cloud are not neighbours in space the launch runs on a Morton-sorted copy (the
    per-point results do not depend on the order of the points) and rows / votes are returned to the cloud's own order."""
    f = len(frames.frame_ids)
    n = xyz_dev.shape[0]
    rles, mask_query, offsets = mask_layout(per_frame, queries, 0, f)
    perm = None
    if REORDER != "never" and n >= 32 and n < (1 << 31) and (REORDER == "always" or (n >= REORDER_MIN_POINTS and float(n) * f >= REORDER_MIN_POINT_VIEWS
                                                                                   and point_order_spread(xyz_dev) > REORDER_SPREAD)):
        perm = morton_order(xyz_dev)
        xyz_dev = xyz_dev[perm].contiguous()                 # position j now holds the cloud's point perm[j]
    out = lift_frames(xyz_dev, intrinsics, frames, rles, mask_query, offsets, len(queries), height, width,
                      want_viewed, depth_thresh, frames_per_chunk)
    rows, mc, vc = out.rows, out.masked_counts, out.viewed_counts
    if perm is not None:
        inv = torch.empty_like(perm)
        inv[perm] = torch.arange(n, device=perm.device)      # the cloud's point p was lifted at position inv[p]
        rows = ops.rows_permute(rows, inv.to(torch.int32), n)
        mc = mc[:, inv].contiguous()
        vc = vc[inv].contiguous() if vc is not None else None
    return LiftedScene(n, split_rows(rows, per_frame, queries), mc, vc if want_viewed else None, f)
