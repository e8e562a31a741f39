"""The reference's function surface ("Python functions"), same names, argument meaning and
return types, executed by libbff.so on the GPU. A notebook or a `_single` script that did
`from projection_2d_to_3d import compute_projected_pts_tensor, ...` keeps working against tools/ of this repo.

numpy in / numpy out for the three lifting functions, torch for the rest, exactly like the reference
(tools/projection_2d_to_3d.py:37,51,73,100,149,277; tools/refinement.py:26,69). Every call goes to the CUDA
library; there is no CPU path.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch

from . import aggregate as _agg, ops
from .rle import runs_for_device

_IDENTITY = np.eye(4, dtype=np.float64)[None]


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("bff_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _lift_one_frame(pts, cam_intr, depth_im, depth_thresh, masks=None):
    """Camera-space points through bff_lift_scene with an identity pose (1*x + 0*y + 0*z + 0 is exact)."""
    dev = _device()
    pts = np.ascontiguousarray(pts, dtype=np.float64)
    if depth_im is None:
        h = w = 1
        depth = torch.zeros((1, 1, 1), dtype=torch.float32, device=dev)
    else:
        h, w = depth_im.shape
        depth = torch.from_numpy(np.ascontiguousarray(depth_im, dtype=np.float32)).to(dev).reshape(1, h, w)
    kw = {}
    if masks is not None and masks.shape[0] > 0:
        m = torch.from_numpy(np.ascontiguousarray(masks != 0)).to(dev).reshape(masks.shape[0], -1)
        kw = dict(masks2d=ops.pack_bits(m), mask_offset=torch.tensor([0, masks.shape[0]], dtype=torch.int32, device=dev))
    return ops.lift_scene(torch.from_numpy(pts).to(dev), torch.from_numpy(_IDENTITY).to(dev), np.asarray(cam_intr, np.float64),
                          depth, h, w, depth_thresh=float(depth_thresh), debug=True, want_counts=False, want_viewed=False, **kw)


def compute_projected_pts_tensor(pts, cam_intr):
    """(N,3) camera-space points, (3,3) intrinsics -> (N,2) int64 [x, y] (projection_2d_to_3d.py:37-48)."""
    out = _lift_one_frame(pts, cam_intr, None, 0.0)
    return out.debug_pix[0].cpu().numpy()


def compute_visibility_mask_tensor(pts, projected_pts, depth_im, depth_thresh=0.005):
    """(N,3) camera-space points, their (N,2) int64 pixels, (H,W) depth -> (N,) bool
    (projection_2d_to_3d.py:51-70): in bounds, depth != 0, |z - depth| < thresh (float64, strict)."""
    from . import _cabi
    dev = _device()
    z = torch.from_numpy(np.ascontiguousarray(np.asarray(pts, dtype=np.float64)[:, 2])).to(dev)
    pix = torch.from_numpy(np.ascontiguousarray(projected_pts, dtype=np.int64)).to(dev)
    h, w = depth_im.shape
    depth = torch.from_numpy(np.ascontiguousarray(depth_im, dtype=np.float32)).to(dev)
    n = z.shape[0]
    vis = torch.zeros((1, ops.words_for(n)), dtype=torch.int32, device=dev)
    _cabi.check(_cabi.lib().bff_visibility(_cabi.ptr(pix), _cabi.ptr(z), n, _cabi.ptr(depth), h, w, float(depth_thresh),
                                           _cabi.ptr(vis), _cabi.current_stream()), "bff_visibility")
    _cabi.launch_count += 1 if n else 0
    return ops.unpack_bits(vis, n)[0].cpu().numpy()


def compute_visible_masked_pts_tensor(scene_pts, projected_pts, visibility_mask, pred_masks):
    """-> (M,N) bool: pred_masks[m, y, x] at the visible points (projection_2d_to_3d.py:73-92)."""
    from . import _cabi
    dev = _device()
    pix = torch.from_numpy(np.ascontiguousarray(projected_pts, dtype=np.int64)).to(dev)
    n = pix.shape[0]
    vis = ops.pack_bits(torch.from_numpy(np.ascontiguousarray(visibility_mask, dtype=bool)).to(dev).reshape(1, n))
    pred_masks = np.asarray(pred_masks)
    m, h, w = pred_masks.shape
    rows = torch.zeros((m, ops.words_for(n)), dtype=torch.int32, device=dev)
    if m and n:
        masks2d = ops.pack_bits(torch.from_numpy(np.ascontiguousarray(pred_masks != 0)).to(dev).reshape(m, -1))
        _cabi.check(_cabi.lib().bff_mask_gather(_cabi.ptr(pix), _cabi.ptr(vis), n, _cabi.ptr(masks2d), m, h, w, _cabi.ptr(rows),
                                                _cabi.current_stream()), "bff_mask_gather")
        _cabi.launch_count += 1
    return ops.unpack_bits(rows, n).cpu().numpy()


def lift_frame(pts, cam_intr, depth_im, pred_masks, depth_thresh=0.08):
    """The three calls above as the kernel actually runs them (one fused pass over one frame):
    returns (projected_pts (N,2) int64, visibility (N,) bool, masked_pts (M,N) bool)."""
    out = _lift_one_frame(pts, cam_intr, depth_im, depth_thresh, np.asarray(pred_masks))
    n = np.asarray(pts).shape[0]
    vis = ops.unpack_bits(out.debug_vis, n)[0].cpu().numpy()
    rows = ops.unpack_bits(out.rows, n).cpu().numpy() if out.rows.shape[0] else np.zeros((0, n), bool)
    return out.debug_pix[0].cpu().numpy(), vis, rows


def calculate_iou(ins_masks: torch.Tensor) -> torch.Tensor:
    """(Ins,N) bool -> (Ins,Ins) float32 (projection_2d_to_3d.py:149-166), on the input's device."""
    dev = ins_masks.device if ins_masks.is_cuda else _device()
    _, _, _, iou = ops.mask_iou(ops.pack_bits(ins_masks.to(dev).to(torch.bool)))
    return iou if ins_masks.is_cuda else iou.cpu()


def calculate_iou_between_stages(mask_1: torch.Tensor, mask_2: torch.Tensor) -> torch.Tensor:
    """(n,N), (m,N) -> (m,n) float32 (refinement.py:69-90)."""
    dev = mask_1.device if mask_1.is_cuda else _device()
    a, b = ops.pack_bits(mask_1.to(dev) != 0), ops.pack_bits(mask_2.to(dev) != 0)
    _, _, _, iou = ops.mask_iou(a, b)
    return iou.T if mask_1.is_cuda else iou.cpu().T


def find_unconnected_subgraphs_tensor(adj_matrix: torch.Tensor) -> List[List[int]]:
    """projection_2d_to_3d.py:250-274 (same lists, without the O(Ins^4) closure)."""
    return _agg.connected_groups(adj_matrix != 0)


def solve_overlapping(ins_masks: torch.Tensor, mask_indeces_to_be_merged: Sequence[Sequence[int]]) -> torch.Tensor:
    """projection_2d_to_3d.py:277-301, in place on packed rows; returns the (Ins,N) bool masks."""
    dev = ins_masks.device if ins_masks.is_cuda else _device()
    rows = ops.pack_bits(ins_masks.to(dev).to(torch.bool))
    inter, _, _, _ = ops.mask_iou(rows, want_iou=False)
    ops.overlap_resolve(rows, _agg.overlap_program(inter, [len(g) for g in mask_indeces_to_be_merged]))
    out = ops.unpack_bits(rows, ins_masks.shape[1])
    return out if ins_masks.is_cuda else out.cpu()


def rle_decode(rle: dict) -> np.ndarray:
    """refinement.py:26-39 -> uint8 (length,) mask; decoded on the device without a per-run Python loop."""
    runs, length = runs_for_device([rle])
    bits = ops.rle_to_bits(runs, 1, length, _device())
    return ops.unpack_bits(bits, length)[0].cpu().numpy().astype(np.uint8)


class _Cfg:
    """Stand-in for the module-level `cfg` the reference's merge_masks reads (projection_2d_to_3d.py:203)."""
    min_aggragated_masks = 2


cfg = _Cfg()


def aggregate(backprojected_3d_masks: dict, iou_threshold=0.25, feature_similarity_threshold=0.75, min_aggragated_masks=None):
    """projection_2d_to_3d.py:100-146: {ins (Ins,N) bool, conf (Ins,), final_class [str]} ->
    ({ins (G,N) bool on the GPU, conf (G,) CPU float32, final_class [str]}, mask_indeces_to_be_merged).
    `feature_similarity_threshold` is accepted and ignored like in the reference."""
    ins = backprojected_3d_masks["ins"]
    dev = ins.device if ins.is_cuda else _device()
    rows = ops.pack_bits(ins.to(dev).to(torch.bool))
    mm = cfg.min_aggragated_masks if min_aggragated_masks is None else min_aggragated_masks
    conf = torch.as_tensor(backprojected_3d_masks["conf"]).cpu()
    agg, conf_agg, labels_agg, groups = _agg.aggregate_rows(rows, conf, list(backprojected_3d_masks["final_class"]),
                                                            iou_threshold, mm)
    if agg is None:
        return {"ins": torch.zeros((0, ins.shape[1]), dtype=torch.bool, device=dev), "conf": torch.zeros(0), "final_class": []}, groups
    return {"ins": ops.unpack_bits(agg, ins.shape[1]), "conf": conf_agg, "final_class": labels_agg}, groups
