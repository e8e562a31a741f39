"""Device operators: torch CUDA tensors in, torch CUDA tensors out, all work done by libbff.so.

Packed masks are int32 tensors of shape (rows, ceil(N/32)) (bit b of word w = element 32*w+b).
torch is used for device memory and streams only; every function raises if the CUDA library
is unavailable (no fallback).
"""
from __future__ import annotations

from ctypes import c_void_p
from typing import Optional, Sequence

import numpy as np
import torch

from . import _cabi
from ._cabi import LiftDesc, check, current_stream, lib, ptr


def words_for(n: int) -> int:
    return (int(n) + 31) // 32


def _cuda(t: torch.Tensor, dtype, what: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _cabi.BffError(f"{what} must be a CUDA tensor (bff_b200 has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{what} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _count(n: int = 1):
    _cabi.launch_count += n


# ----------------------------------------------------------------------------- lifting
class LiftOutput:
    __slots__ = ("rows", "masked_counts", "viewed_counts", "debug_pix", "debug_vis")

    def __init__(self, rows, mc, vc, pix=None, vis=None):
        self.rows, self.masked_counts, self.viewed_counts, self.debug_pix, self.debug_vis = rows, mc, vc, pix, vis


def lift_scene(xyz: torch.Tensor, pose_inv: torch.Tensor, intrinsics, depth: torch.Tensor, height: int, width: int,
               masks2d: Optional[torch.Tensor] = None, mask_offset: Optional[torch.Tensor] = None,
               mask_query: Optional[torch.Tensor] = None, n_queries: int = 1,
               frame_viewed: Optional[torch.Tensor] = None, depth_thresh: float = 0.08,
               depth_scale: float = 1000.0, depth_range: Optional[Sequence[float]] = None,
               want_rows: bool = True, want_counts: bool = True, want_viewed: bool = True,
               debug: bool = False, frames_per_chunk: int = 0, timing_events=None,
               native_size: Optional[Sequence[int]] = None, row_slices: Optional[Sequence[int]] = None,
               votes: Optional[torch.Tensor] = None, peer_rows: Optional[Sequence[int]] = None, peer_row0: int = 0) -> LiftOutput:
    """One launch over all frames of a scene (bff_lift_scene).

    xyz         (N,4) float32 [x,y,z,_] or (N,3) float64, CUDA
    pose_inv    (F,4,4) float64 CUDA = np.linalg.inv(cam_pose) per frame (projection_2d_to_3d.py:425)
    intrinsics  (3,3) float64 HOST (numpy / CPU tensor): passed to the kernel by value
    depth       (F,H,W) float32 metres or uint16/int16-storage raw sensor units, CUDA; with
                native_size=(Hs,Ws) != (H,W): (F,Hs,Ws) raw sensor units sampled in the kernel exactly like
                cv2.resize(depth, (W,H)) (projection_2d_to_3d.py:436) -- BFF_DEPTH_U16_BILINEAR
    masks2d     (sumM, ceil(H*W/32)) int32 packed 2D masks in frame order; mask_offset (F+1,) int32
    row_slices  (G, slice_words): rows come back as (G, sumM, slice_words) -- the words of every row cut into G column
                slices, slice-major (the send layout of the point-sharded multi-GPU exchange, dist.py)
    votes       optional (Q+1, N) int32 CUDA buffer that receives [masked_counts | viewed_counts] (one all-reduce)
    peer_rows   with row_slices: G device addresses (ints) of the ranks' receive buffers; slice p of the rows is stored
                straight into buffer p at row peer_row0 + m (fused exchange over peer memory); no rows tensor is returned
    Returns packed rows (sumM, Wd) int32, masked_counts (Q,N) int32, viewed_counts (N,) int32."""
    dev = xyz.device
    if xyz.dtype == torch.float32:
        if xyz.dim() != 2 or xyz.shape[1] != 4:
            raise ValueError("float32 xyz must be (N,4)")
        kind = _cabi.BFF_XYZ_F32X4
    elif xyz.dtype == torch.float64:
        if xyz.dim() != 2 or xyz.shape[1] != 3:
            raise ValueError("float64 xyz must be (N,3)")
        kind = _cabi.BFF_XYZ_F64X3
    else:
        raise TypeError("xyz must be float32 (N,4) or float64 (N,3)")
    xyz = _cuda(xyz, xyz.dtype, "xyz")
    n = xyz.shape[0]
    pose_inv = _cuda(pose_inv, torch.float64, "pose_inv").reshape(-1, 4, 4)
    f = pose_inv.shape[0]
    k_host = np.ascontiguousarray(np.asarray(intrinsics, dtype=np.float64).reshape(3, 3))
    if depth.dtype == torch.float32:
        dkind = _cabi.BFF_DEPTH_F32
    elif depth.dtype in (torch.int16, torch.uint16):
        dkind = _cabi.BFF_DEPTH_U16
    else:
        raise TypeError("depth must be float32 or 16-bit raw")
    depth = _cuda(depth, depth.dtype, "depth")
    taps = None
    if native_size is not None and (int(native_size[0]), int(native_size[1])) != (int(height), int(width)):
        if dkind != _cabi.BFF_DEPTH_U16:
            raise TypeError("in-kernel resize needs raw 16-bit depth")
        dkind = _cabi.BFF_DEPTH_U16_BILINEAR
        hs, ws = int(native_size[0]), int(native_size[1])
        if depth.numel() != f * hs * ws:
            raise ValueError(f"depth has {depth.numel()} elements, expected {f}x{hs}x{ws}")
        taps = (resize_table(width, ws, dev), resize_table(height, hs, dev))
    elif depth.numel() != f * height * width:
        raise ValueError(f"depth has {depth.numel()} elements, expected {f}x{height}x{width}")
    n_masks = 0
    if masks2d is not None and masks2d.shape[0] > 0:
        masks2d = _cuda(masks2d, torch.int32, "masks2d")
        n_masks = masks2d.shape[0]
        if masks2d.shape[1] != words_for(height * width):
            raise ValueError("masks2d has the wrong number of words per mask")
        mask_offset = _cuda(mask_offset, torch.int32, "mask_offset")
        if mask_offset.numel() != f + 1:
            raise ValueError("mask_offset must have F+1 entries")
        if mask_query is not None:
            mask_query = _cuda(mask_query, torch.int32, "mask_query")
    else:
        masks2d = mask_offset = mask_query = None
    if frame_viewed is not None:
        frame_viewed = _cuda(frame_viewed, torch.uint8, "frame_viewed")
    wd = words_for(n)
    if row_slices is not None:
        g_sl, sw = int(row_slices[0]), int(row_slices[1])
        if g_sl * sw < wd:
            raise ValueError("row slices do not cover the row")
        rows = torch.empty((g_sl, n_masks, sw), dtype=torch.int32, device=dev) if (want_rows and peer_rows is None) else None
    else:
        rows = torch.empty((n_masks, wd), dtype=torch.int32, device=dev) if want_rows else None
    if votes is not None:
        if votes.dtype != torch.int32 or tuple(votes.shape) != (n_queries + 1, n) or not votes.is_contiguous() or not votes.is_cuda:
            raise ValueError("votes must be a contiguous (Q+1, N) int32 CUDA tensor")
        mc, vc = votes[:n_queries], votes[n_queries]
        if not want_viewed:
            vc.zero_()
    else:
        mc = torch.empty((n_queries, n), dtype=torch.int32, device=dev) if want_counts else None
        vc = torch.empty((n,), dtype=torch.int32, device=dev) if want_viewed else None
    pix = torch.empty((f, n, 2), dtype=torch.int64, device=dev) if debug else None
    vis = torch.empty((f, wd), dtype=torch.int32, device=dev) if debug else None
    d = LiftDesc()
    d.xyz, d.xyz_kind, d.n_points = ptr(xyz), kind, n
    d.pose_inv, d.intrinsics, d.n_frames = ptr(pose_inv), ptr(k_host), f
    d.depth, d.depth_kind, d.depth_scale = ptr(depth), dkind, float(depth_scale)
    d.height, d.width, d.depth_thresh = int(height), int(width), float(depth_thresh)
    if depth_range is None:
        d.depth_min, d.depth_max = 1.0, 0.0
    else:
        d.depth_min, d.depth_max = float(depth_range[0]), float(depth_range[1])
    d.masks2d, d.mask_offset, d.mask_query = ptr(masks2d), ptr(mask_offset), ptr(mask_query)
    d.n_masks, d.n_queries, d.frame_viewed = n_masks, int(n_queries), ptr(frame_viewed)
    d.rows, d.masked_counts, d.viewed_counts = ptr(rows), ptr(mc), ptr(vc if (want_viewed or votes is None) else None)
    d.debug_pix, d.debug_vis = ptr(pix), ptr(vis)
    if row_slices is not None:
        d.n_row_slices, d.row_slice_words = g_sl, sw
        if peer_rows is not None:
            if len(peer_rows) != g_sl or g_sl > 8:
                raise ValueError("peer_rows needs one address per slice (at most 8)")
            for i_p, a_p in enumerate(peer_rows):
                d.peer_rows[i_p] = int(a_p)
            d.peer_row0 = int(peer_row0)
    d.frames_per_chunk = int(frames_per_chunk)
    if taps is not None:
        d.src_height, d.src_width, d.resize_x, d.resize_y = hs, ws, ptr(taps[0]), ptr(taps[1])
    if timing_events is not None:      # two recorded torch.cuda.Event(enable_timing=True): re-recorded around the launch
        d.ev_start, d.ev_stop = c_void_p(timing_events[0].cuda_event), c_void_p(timing_events[1].cuda_event)
    check(lib().bff_lift_scene(d, current_stream()), "bff_lift_scene")
    _count(lib().bff_lift_scene_launches(d))
    return LiftOutput(rows, mc, vc, pix, vis)


_tap_cache = {}


def resize_table(n_dst: int, n_src: int, device=None):
    """bff_resize_table: (n_dst, 2) int32 rows [source index, float32 weight bits] of one axis of
    cv2.resize(..., INTER_LINEAR). Host numpy array, or a cached CUDA copy when `device` is given."""
    key = (int(n_dst), int(n_src))
    if key not in _tap_cache:
        t = np.empty((key[0], 2), np.int32)
        check(lib().bff_resize_table(key[0], key[1], ptr(t)), "bff_resize_table")
        _tap_cache[key] = {"host": t}
    ent = _tap_cache[key]
    if device is None:
        return ent["host"]
    dk = str(device)
    if dk not in ent:
        ent[dk] = torch.from_numpy(ent["host"]).to(device)
    return ent[dk]


def depth_resize(raw: torch.Tensor, height: int, width: int, depth_scale: float = 1000.0) -> torch.Tensor:
    """(F,Hs,Ws) raw 16-bit depth on the GPU -> (F,H,W) float32 metres = cv2.resize(raw.astype(f32)/scale, (W,H))
    (projection_2d_to_3d.py:431-436), bff_depth_resize."""
    if raw.dtype not in (torch.int16, torch.uint16) or raw.dim() != 3:
        raise TypeError("raw depth must be a (F,Hs,Ws) 16-bit tensor")
    raw = _cuda(raw, raw.dtype, "raw")
    f, hs, ws = raw.shape
    out = torch.empty((f, int(height), int(width)), dtype=torch.float32, device=raw.device)
    tx, ty = resize_table(width, ws, raw.device), resize_table(height, hs, raw.device)
    check(lib().bff_depth_resize(ptr(raw), f, hs, ws, float(depth_scale), ptr(tx), ptr(ty), int(height), int(width), ptr(out),
                                 current_stream()), "bff_depth_resize")
    if f:
        _count(1)
    return out


# ----------------------------------------------------------------------------- mask algebra
def mask_iou(a: torch.Tensor, b: Optional[torch.Tensor] = None, want_iou: bool = True):
    """Intersections, row popcounts and (optionally) float32 IoU of packed rows.
    b=None computes the square matrix of `a` (calculate_iou). Returns (inter, pop_a, pop_b, iou)."""
    a = _cuda(a, torch.int32, "a")
    same = b is None
    b = a if same else _cuda(b, torch.int32, "b")
    if a.shape[1] != b.shape[1]:
        raise ValueError("row width mismatch")
    na, nb, wd = a.shape[0], b.shape[0], a.shape[1]
    dev = a.device
    inter = torch.zeros((na, nb), dtype=torch.int32, device=dev)
    pa = torch.zeros((na,), dtype=torch.int32, device=dev)
    pb = pa if same else torch.zeros((nb,), dtype=torch.int32, device=dev)
    iou = torch.empty((na, nb), dtype=torch.float32, device=dev) if want_iou else None
    check(lib().bff_mask_iou(ptr(a), na, ptr(b), nb, wd, ptr(inter), ptr(pa), ptr(pb), ptr(iou), current_stream()),
          "bff_mask_iou")
    if na and nb:
        _count((1 if same else 2) + 1 + (1 if want_iou else 0))
    return inter, pa, pb, iou


def rows_or_reduce(rows: torch.Tensor, groups: Sequence[Sequence[int]]) -> torch.Tensor:
    """out[g] = OR of rows[k] for k in groups[g]."""
    rows = _cuda(rows, torch.int32, "rows")
    wd = rows.shape[1]
    idx = np.asarray([k for g in groups for k in g], dtype=np.int32)
    seg = np.zeros(len(groups) + 1, dtype=np.int32)
    seg[1:] = np.cumsum([len(g) for g in groups])
    if idx.size and (idx.min() < 0 or idx.max() >= rows.shape[0]):
        raise IndexError("group member out of range")
    out = torch.zeros((len(groups), wd), dtype=torch.int32, device=rows.device)
    if len(groups) and wd:
        d_idx = torch.from_numpy(idx).to(rows.device)
        d_seg = torch.from_numpy(seg).to(rows.device)
        check(lib().bff_rows_or_reduce(ptr(rows), wd, ptr(d_idx), ptr(d_seg), len(groups), ptr(out), current_stream()),
              "bff_rows_or_reduce")
        _count()
    return out


def counts_max(mc: torch.Tensor, vc: Optional[torch.Tensor]):
    mc = _cuda(mc, torch.int32, "mc").reshape(-1)
    if vc is not None:
        vc = _cuda(vc, torch.int32, "vc").reshape(-1)
    out = torch.zeros(2, dtype=torch.int32, device=mc.device)
    check(lib().bff_counts_max(ptr(mc), ptr(vc), mc.numel(), ptr(out), current_stream()), "bff_counts_max")
    _count()
    m = out.cpu()
    return int(m[0]), int(m[1])


def vote_hist(mc: torch.Tensor, vc: Optional[torch.Tensor], mc_bins: int, vc_bins: int) -> torch.Tensor:
    mc = _cuda(mc, torch.int32, "mc").reshape(-1)
    if vc is not None:
        vc = _cuda(vc, torch.int32, "vc").reshape(-1)
    hist = torch.empty((mc_bins, vc_bins), dtype=torch.int32, device=mc.device)
    check(lib().bff_vote_hist(ptr(mc), ptr(vc), mc.numel(), mc_bins, vc_bins, ptr(hist), current_stream()), "bff_vote_hist")
    _count()
    return hist


def vote_unique(hist: torch.Tensor, mode: int) -> torch.Tensor:
    """Distinct values of the non-empty histogram bins, sorted: what masked_counts.unique() (mode 1) or
    detected_ratio.unique() (mode 0, f32(mc) / (f32(vc) + 1)) returns in the reference (projection_2d_to_3d.py:512-517, :571-575)."""
    hist = _cuda(hist, torch.int32, "hist")
    mb, vb = hist.shape
    tsize = int(lib().bff_vote_unique_table_size(mb, vb))
    table = torch.empty((tsize,), dtype=torch.int32, device=hist.device)
    vals = torch.empty((mb * vb,), dtype=torch.float32, device=hist.device)
    cnt = torch.zeros((1,), dtype=torch.int32, device=hist.device)
    check(lib().bff_vote_unique(ptr(hist), mb, vb, int(mode), ptr(table), tsize, ptr(vals), ptr(cnt), current_stream()), "bff_vote_unique")
    _count(2)
    return torch.sort(vals[: int(cnt.item())]).values


def vote_unique_static(hist: torch.Tensor, mode: int, values: torch.Tensor, count: torch.Tensor) -> None:
    """bff_vote_unique_static: vote_unique with the scratch table inside the library and no allocation anywhere -- the
    caller brings `values` (float32, >= hist.numel()) and `count` (int32, 1) and reads them itself (nothing is allocated,
    copied or synchronised here). At most _cabi.VOTE_UNIQUE_STATIC_MAX_BINS histogram bins; calls must not overlap on the GPU."""
    if not (hist.is_cuda and hist.dtype == torch.int32 and hist.is_contiguous() and hist.dim() == 2):
        raise TypeError("hist must be a contiguous 2-D int32 CUDA tensor")
    if not (values.is_cuda and values.dtype == torch.float32 and values.is_contiguous() and values.numel() >= hist.numel()):
        raise TypeError("values must be a contiguous float32 CUDA tensor with one slot per histogram bin")
    if not (count.is_cuda and count.dtype == torch.int32 and count.numel() >= 1):
        raise TypeError("count must be an int32 CUDA tensor")
    check(lib().bff_vote_unique_static(ptr(hist), hist.shape[0], hist.shape[1], int(mode), ptr(values), ptr(count), current_stream()),
          "bff_vote_unique_static")
    _count(2)


def point_keep(mc: torch.Tensor, vc: Optional[torch.Tensor], mode: int, thr: float) -> torch.Tensor:
    mc = _cuda(mc, torch.int32, "mc").reshape(-1)
    if vc is not None:
        vc = _cuda(vc, torch.int32, "vc").reshape(-1)
    keep = torch.empty((words_for(mc.numel()),), dtype=torch.int32, device=mc.device)
    check(lib().bff_point_keep(ptr(mc), ptr(vc), mc.numel(), int(mode), float(thr), ptr(keep), current_stream()),
          "bff_point_keep")
    _count()
    return keep


def _row_program(fn_name: str, rows: torch.Tensor, program: Sequence[Sequence[int]]) -> torch.Tensor:
    if not rows.is_contiguous():
        raise ValueError("in-place row programs need a contiguous rows tensor")
    rows = _cuda(rows, torch.int32, "rows")
    prog = np.asarray(program, dtype=np.int32).reshape(-1, 2)
    if prog.shape[0] and rows.shape[0] and rows.shape[1]:
        if prog.min() < 0 or prog.max() >= rows.shape[0]:
            raise IndexError("program row out of range")
        d_prog = torch.from_numpy(prog).to(rows.device)
        check(getattr(lib(), fn_name)(ptr(rows), rows.shape[0], rows.shape[1], ptr(d_prog), prog.shape[0], current_stream()),
              fn_name)
        _count()
    return rows


def overlap_resolve(rows: torch.Tensor, program: Sequence[Sequence[int]]) -> torch.Tensor:
    """In place, sequentially: for (dst, src) in program: rows[dst] &= ~rows[src]."""
    return _row_program("bff_overlap_resolve", rows, program)


def rows_or_into(rows: torch.Tensor, program: Sequence[Sequence[int]]) -> torch.Tensor:
    """In place, sequentially: for (dst, src) in program: rows[dst] |= rows[src]."""
    return _row_program("bff_rows_or_into", rows, program)


def rows_and_popcount(rows: torch.Tensor, keep: Optional[torch.Tensor] = None) -> torch.Tensor:
    """rows &= keep (in place, if given); returns int32 popcounts per row."""
    rows = _cuda(rows, torch.int32, "rows")
    if keep is not None:
        keep = _cuda(keep, torch.int32, "keep")
        if keep.numel() != rows.shape[1]:
            raise ValueError("keep width mismatch")
    pop = torch.zeros((rows.shape[0],), dtype=torch.int32, device=rows.device)
    check(lib().bff_rows_and_popcount(ptr(rows), rows.shape[0], rows.shape[1], ptr(keep), ptr(pop), current_stream()),
          "bff_rows_and_popcount")
    if rows.shape[0] and rows.shape[1]:
        _count()
    return pop


def pack_bits(dense: torch.Tensor) -> torch.Tensor:
    """(R,N) bool/uint8 CUDA -> (R, ceil(N/32)) int32."""
    if dense.dtype == torch.bool:
        dense = dense.view(torch.uint8)
    dense = _cuda(dense, torch.uint8, "dense")
    r, n = dense.shape
    out = torch.zeros((r, words_for(n)), dtype=torch.int32, device=dense.device)
    check(lib().bff_pack_bits(ptr(dense), r, n, ptr(out), current_stream()), "bff_pack_bits")
    if r and n:
        _count()
    return out


def unpack_bits(packed: torch.Tensor, n: int) -> torch.Tensor:
    """(R, Wd) int32 -> (R,N) bool CUDA."""
    packed = _cuda(packed, torch.int32, "packed")
    r = packed.shape[0]
    if packed.shape[1] != words_for(n):
        raise ValueError("packed width does not match n")
    out = torch.empty((r, n), dtype=torch.uint8, device=packed.device)
    check(lib().bff_unpack_bits(ptr(packed), r, n, ptr(out), current_stream()), "bff_unpack_bits")
    if r and n:
        _count()
    return out.view(torch.bool)


def rle_to_bits_dev(runs: torch.Tensor, n_rows: int, length: int) -> torch.Tensor:
    """runs (R,3) int32 CUDA [row, lo, hi) -> packed (n_rows, ceil(length/32)) int32."""
    runs = _cuda(runs, torch.int32, "runs").reshape(-1, 3)
    wd = words_for(length)
    out = torch.empty((n_rows, wd), dtype=torch.int32, device=runs.device)
    check(lib().bff_rle_to_bits(ptr(runs) if runs.shape[0] else None, runs.shape[0], n_rows, wd, ptr(out),
                                current_stream()), "bff_rle_to_bits")
    if n_rows and wd and runs.shape[0]:
        _count()
    return out


def rle_pairs_to_bits_dev(counts: torch.Tensor, pair_offset: torch.Tensor, length: int) -> torch.Tensor:
    """RLE wire format -> packed rows on the device (bff_rle_pairs_to_bits): counts int64 CUDA = the masks' (1-based start,
    length) pairs back to back, pair_offset (R+1,) int64 CUDA. The reference's int32 / slice arithmetic runs in the kernel."""
    counts = _cuda(counts, torch.int64, "counts").reshape(-1)
    pair_offset = _cuda(pair_offset, torch.int64, "pair_offset").reshape(-1)
    n_rows = pair_offset.numel() - 1
    out = torch.empty((n_rows, words_for(length)), dtype=torch.int32, device=counts.device)
    check(lib().bff_rle_pairs_to_bits(ptr(counts) if counts.numel() else None, ptr(pair_offset), n_rows, counts.numel() // 2, int(length), ptr(out),
                                      current_stream()), "bff_rle_pairs_to_bits")
    if n_rows and length and counts.numel():
        _count()
    return out


def rle_to_bits(runs: np.ndarray, n_rows: int, length: int, device) -> torch.Tensor:
    """runs (R,3) int32 host [row, lo, hi) -> packed (n_rows, ceil(length/32)) int32 on `device`."""
    runs = np.ascontiguousarray(runs, dtype=np.int32).reshape(-1, 3)
    return rle_to_bits_dev(torch.from_numpy(runs).to(device), n_rows, length)


def to_host(t: torch.Tensor) -> torch.Tensor:
    """Device -> host through pinned memory (torch's caching host allocator recycles the blocks), about 3x
    the bandwidth of `.cpu()` into pageable memory for the multi-megabyte (Ins, N) bool outputs."""
    if not t.is_cuda or t.numel() == 0:
        return t.cpu()
    out = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    out.copy_(t, non_blocking=True)
    torch.cuda.current_stream(t.device).synchronize()
    return out


def upload(dst: torch.Tensor, src_pinned: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
    """dst (CUDA) <- src (pinned host) through bff_upload: an SM-driven copy that leaves the copy engines free.
    Enqueued on `stream` (default: current)."""
    if not (dst.is_cuda and dst.is_contiguous() and src_pinned.is_contiguous()):
        raise ValueError("upload needs a contiguous CUDA destination and a contiguous host source")
    if not src_pinned.is_pinned():
        raise ValueError("upload needs pinned host memory")
    nbytes = src_pinned.numel() * src_pinned.element_size()
    if nbytes != dst.numel() * dst.element_size():
        raise ValueError("size mismatch")
    st = stream if stream is not None else torch.cuda.current_stream(dst.device)
    check(lib().bff_upload(ptr(dst), c_void_p(src_pinned.data_ptr()), nbytes, c_void_p(st.cuda_stream)), "bff_upload")
    if nbytes:
        _count()
    return dst


_STAGE_BYTES = 1 << 20
_stage_rings: dict = {}


def small_to_device(t: torch.Tensor, device) -> torch.Tensor:
    """Small CPU tensor -> CUDA tensor through a pinned staging ring and bff_upload (an SM copy kernel on the current
    stream) instead of a copy-engine transfer: index vectors of a few bytes must not queue behind a scene-sized
    cudaMemcpyAsync that another stream has put on the H2D engine. Falls back to `.to(device)` for large tensors."""
    import threading
    device = torch.device(device)
    t = t.contiguous()
    nbytes = t.numel() * t.element_size()
    if nbytes == 0 or nbytes > _STAGE_BYTES // 8 or device.type != "cuda":
        return t.to(device)
    key = (device.index or 0, threading.get_ident())
    ring = _stage_rings.get(key)
    if ring is None:
        ring = _stage_rings[key] = [torch.empty(_STAGE_BYTES, dtype=torch.uint8).pin_memory(), 0]
    need = (nbytes + 255) & ~255
    if ring[1] + need > _STAGE_BYTES:
        # wrap (once per ~4000 index vectors): an upload kernel that has not run yet may still read the oldest slots
        torch.cuda.synchronize(device)
        ring[1] = 0
    slot = ring[0][ring[1]:ring[1] + nbytes]
    ring[1] += need
    slot.copy_(t.reshape(-1).view(torch.uint8))
    out = torch.empty(t.shape, dtype=t.dtype, device=device)
    check(lib().bff_upload(ptr(out), c_void_p(slot.data_ptr()), nbytes, current_stream()), "bff_upload")
    _count()
    return out


def labels_to_rows(labels: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """labels (N,) int64 CUDA, ids (G,) int64 CUDA -> (G, ceil(N/32)) int32 packed rows, bit p of row g = labels[p] == ids[g]
    (bff_labels_to_rows; evaluator side, scannetv2_inst_eval.py:348)."""
    labels = _cuda(labels, torch.int64, "labels").reshape(-1)
    ids = _cuda(ids, torch.int64, "ids").reshape(-1)
    n, g = labels.shape[0], ids.shape[0]
    out = torch.zeros((g, words_for(n)), dtype=torch.int32, device=labels.device)
    check(lib().bff_labels_to_rows(ptr(labels), n, ptr(ids), g, ptr(out), current_stream()), "bff_labels_to_rows")
    if n and g:
        _count()
    return out


def rows_gather(rows: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """rows[idx] for packed rows with the int32 index vector already on the device (bff_rows_gather_dev): one kernel of
    this library, no ATen indexing kernel and no synchronisation."""
    rows = _cuda(rows, torch.int32, "rows")
    idx = _cuda(idx, torch.int32, "idx").reshape(-1)
    out = torch.empty((idx.numel(), rows.shape[1]), dtype=torch.int32, device=rows.device)
    check(lib().bff_rows_gather_dev(ptr(rows), rows.shape[1], ptr(idx), idx.numel(), ptr(out), current_stream()), "bff_rows_gather_dev")
    if idx.numel() and rows.shape[1]:
        _count()
    return out


def rows_slice(rows: torch.Tensor, word0: int, slice_words: int) -> torch.Tensor:
    """Words [word0, word0 + slice_words) of every packed row, zero beyond the row (bff_rows_slice)."""
    rows = _cuda(rows, torch.int32, "rows")
    out = torch.empty((rows.shape[0], int(slice_words)), dtype=torch.int32, device=rows.device)
    check(lib().bff_rows_slice(ptr(rows), rows.shape[0], rows.shape[1], int(word0), int(slice_words), ptr(out), current_stream()), "bff_rows_slice")
    if rows.shape[0] and slice_words:
        _count()
    return out


def rows_unslice(slices: torch.Tensor, words: int) -> torch.Tensor:
    """(G, R, slice_words) column slices -> (R, words) rows (bff_rows_unslice)."""
    slices = _cuda(slices, torch.int32, "slices")
    g, r, sw = slices.shape
    out = torch.empty((r, int(words)), dtype=torch.int32, device=slices.device)
    check(lib().bff_rows_unslice(ptr(slices), g, r, sw, int(words), ptr(out), current_stream()), "bff_rows_unslice")
    if r and words:
        _count()
    return out


def rows_permute(rows: torch.Tensor, src_index: torch.Tensor, n_points: int) -> torch.Tensor:
    """out bit p of every row = rows bit src_index[p] (bff_rows_permute). rows (R,Wd) int32, src_index (N,) int32, CUDA."""
    rows = _cuda(rows, torch.int32, "rows")
    src_index = _cuda(src_index, torch.int32, "src_index").reshape(-1)
    if src_index.numel() != n_points or rows.shape[1] != words_for(n_points):
        raise ValueError("rows_permute: size mismatch")
    out = torch.empty_like(rows)
    check(lib().bff_rows_permute(ptr(rows), rows.shape[0], int(n_points), ptr(src_index), ptr(out), current_stream()), "bff_rows_permute")
    if rows.shape[0] and n_points:
        _count()
    return out
