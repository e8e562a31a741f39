"""Aggregation, point filtering and overlap resolution of lifted 3D masks on PACKED rows.

Host-side mirror of aggregate / calculate_iou / calculate_feature_similarity / merge_masks /
find_unconnected_subgraphs_tensor / solve_overlapping and the filtering block of
tools/projection_2d_to_3d.py (:100-301, :512-623). The O(Ins^2 N) and O(Ins N) work runs in
libbff.so on bit-packed rows; what stays on the host are decisions over tiny matrices, made
with the same torch CPU ops as the reference so comparison/promotion semantics are identical.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _log, ops
from ._log import log


def label_equality(labels: Sequence[str]) -> torch.Tensor:
    """calculate_feature_similarity (projection_2d_to_3d.py:169-187): exact string equality.
    (The reference's similarity threshold argument is unused)"""
    _, inv = np.unique(np.asarray(list(labels), dtype=object).astype(str), return_inverse=True) if len(labels) else ((), np.zeros(0, np.int64))
    v = torch.from_numpy(np.asarray(inv, dtype=np.int64))
    return v.unsqueeze(0) == v.unsqueeze(1)


def _connected_groups_general(a: np.ndarray) -> List[List[int]]:
    n = a.shape[0]
    nbrs = [np.flatnonzero(a[i]) for i in range(n)]
    visited = np.zeros(n, dtype=bool)
    out: List[List[int]] = []
    for i in range(n):
        if visited[i]:
            continue
        reach = np.zeros(n, dtype=bool)
        stack = list(nbrs[i])
        reach[stack] = True
        while stack:
            u = stack.pop()
            for v in nbrs[u]:
                if not reach[v]:
                    reach[v] = True
                    stack.append(v)
        out.append([int(k) for k in np.flatnonzero(reach)])
        visited[reach] = True
    return out


def connected_groups(adj: torch.Tensor) -> List[List[int]]:
    """find_unconnected_subgraphs_tensor (projection_2d_to_3d.py:250-274) without the O(Ins^4)
    closure: for every not-yet-visited node in index order, the ascending list of nodes reachable
    from it by a path of length >= 1. A node without self edge and without neighbours (an empty
    lifted row: its IoU diagonal is NaN) yields [] and marks nothing visited, like the reference.
    The merge matrix is symmetric (IoU and label equality both are), in which case the lists are the
    connected components (scipy csgraph, C speed) ordered by smallest member."""
    a = np.ascontiguousarray(adj.to(torch.bool).cpu().numpy())
    n = a.shape[0]
    if n == 0:
        return []
    if not np.array_equal(a, a.T):
        return _connected_groups_general(a)
    from scipy.sparse import csr_matrix
    from scipy.sparse.csgraph import connected_components
    r, c = np.nonzero(a)
    _, lab = connected_components(csr_matrix((np.ones(r.size, np.int8), (r, c)), shape=(n, n)), directed=False)
    has_edge = np.zeros(n, bool)
    has_edge[r] = True
    order = np.argsort(lab, kind="stable")                    # members ascending inside each component
    bounds = np.flatnonzero(np.diff(lab[order])) + 1
    comps = np.split(order, bounds)
    comps.sort(key=lambda c: c[0])                            # emitted when their smallest node is reached
    return [[int(k) for k in c] if has_edge[c[0]] else [] for c in comps]


def aggregate_rows(rows: torch.Tensor, conf: torch.Tensor, labels: Sequence[str], iou_thres: float,
                   min_members: int):
    """aggregate + merge_masks (projection_2d_to_3d.py:100-146,190-247) on packed rows.
    rows (Ins,Wd) int32 CUDA; conf (Ins,) CPU tensor as torch.tensor(list of 0-d) builds it.
    Returns (agg_rows CUDA or None, conf_agg CPU tensor, labels_agg, groups)."""
    _, _, _, iou = ops.mask_iou(rows)                           # K7: popcount intersections, exact fp32 IoU
    merge = label_equality(labels) & (iou.cpu() > iou_thres)    # python float vs f32 tensor: torch semantics
    groups = [g for g in connected_groups(merge) if len(g) >= min_members]
    log("masks_to_be_merged", groups)
    used = [g for g in groups if g != []]
    if not used:
        return None, None, [], []
    agg = ops.rows_or_reduce(rows, used)                        # K10
    # conf = sum(conf_k) / len, the sum taken sequentially in index order IN THE CONFIDENCES' OWN DTYPE (:220-224: Python's
    # sum() over 0-d tensors, then torch.tensor(list) keeps the dtype: float64 confidences give float64 means)
    c = conf.numpy()
    dt = c.dtype if c.dtype in (np.dtype(np.float32), np.dtype(np.float64)) else np.dtype(np.float32)
    confs = np.empty(len(used), dt)
    for i, g in enumerate(used):
        confs[i] = np.cumsum(c[g].astype(dt, copy=False), dtype=dt)[-1] / dt.type(len(g))
    return agg, torch.from_numpy(confs), [labels[g[0]] for g in used], groups


def keep_points(mc: torch.Tensor, vc: Optional[torch.Tensor], cfg) -> torch.Tensor:
    """Point filter (projection_2d_to_3d.py:512-583) -> packed keep bits (Wd,) int32 CUDA.
    The reference's `tensor.unique()` over N points is replaced by the (masked, viewed) histogram:
    every distinct value of the ratio comes from a non-empty bin, so the unique list is the same."""
    mc = mc.reshape(-1)
    if cfg["if_occurance_threshold"]:
        mmax, _ = ops.counts_max(mc, None)
        hist = ops.vote_hist(mc, None, mmax + 1, 1).cpu().reshape(-1)
        u = torch.nonzero(hist > 0).reshape(-1).to(torch.float32)       # ascending == unique() order
        t = u[math.floor(cfg["occurance_threshold"] * u.shape[0])]
        return ops.point_keep(mc, None, 1, float(t))
    if cfg["if_detected_ratio_threshold"]:
        mmax, vmax = ops.counts_max(mc, vc)
        hist = ops.vote_hist(mc, vc, mmax + 1, vmax + 1).cpu()
        nz = torch.nonzero(hist > 0)
        ratio = nz[:, 0].to(torch.float32) / (nz[:, 1].to(torch.float32) + 1)   # :571 in float32
        u = ratio.unique()
        t = u[math.floor(cfg["detected_ratio_threshold"] * u.shape[0])]
        log("detected_ratio_thres_value", t)
        return ops.point_keep(mc, vc, 0, float(t))
    return ops.point_keep(mc, None, 2, 0.0)


def overlap_program(inter: torch.Tensor, group_sizes: Sequence[int]) -> List[Tuple[int, int]]:
    """solve_overlapping's schedule (projection_2d_to_3d.py:285-299): pairs i<j that overlap in the
    incoming rows, in (i, j) order; the mask merged from more members keeps the shared points."""
    ov = (inter.cpu() > 0).numpy()
    prog = []
    n = ov.shape[0]
    for i in range(n):
        for j in np.flatnonzero(ov[i, i + 1:]) + i + 1:
            j = int(j)
            if group_sizes[i] > group_sizes[j]:
                prog.append((j, i))          # rows[j] &= ~rows[i]
            else:
                prog.append((i, j))          # rows[i] &= ~rows[j]
    return prog


def filter_and_select(agg: torch.Tensor, conf_agg: torch.Tensor, labels_agg: Sequence[str], groups: List[List[int]],
                      keep_bits: torch.Tensor, n_points: int, cfg, to_host: bool = True):
    """projection_2d_to_3d.py:583-634: overlap resolution, AND with the kept points, row filter,
    and (the dict that is saved, the packed rows of its masks)."""
    inter, before, _, _ = ops.mask_iou(agg, want_iou=False)
    sizes = [len(g) for g in groups]          # as the reference: indexed by position in the merge list
    ops.overlap_resolve(agg, overlap_program(inter, sizes))
    after = ops.rows_and_popcount(agg, keep_bits)
    before64, after64 = before.cpu().to(torch.int64), after.cpu().to(torch.int64)
    sel = (after64 > cfg["remove_small_masks"]) & (after64 > cfg["remove_filtered_masks"] * before64)
    idx = torch.nonzero(sel).reshape(-1)
    # every row kept (the usual case): the buffer itself, no index upload and no gather
    out_rows = agg if idx.numel() == agg.shape[0] else agg[ops.small_to_device(idx, agg.device)]
    ins = ops.unpack_bits(out_rows, n_points)
    log("after filtering", ins.shape)
    result = {"ins": ops.to_host(ins) if to_host else ins, "conf": conf_agg[sel], "final_class": [labels_agg[i] for i in idx.tolist()]}
    return result, out_rows


# ----------------------------------------------------------------------------- fused C++ pipeline
class _FusedJob:
    """Inputs, descriptor and host output buffers of one bff_stage2_aggregate job (kept alive until the call returns)."""

    def __init__(self, rows, conf, labels, mc, vc, n_points, cfg, pairs=None):
        from . import _cabi
        n = rows.shape[0]
        self.labels, self.n = labels, n
        if n and isinstance(labels, (list, tuple)) and labels.count(labels[0]) == n:   # one phrase for every mask (the usual case: labels = [query] * M)
            label_id = np.zeros(n, dtype=np.int32)
        else:
            ids: dict = {}
            label_id = np.asarray([ids.setdefault(l, len(ids)) for l in labels], dtype=np.int32)
        conf32 = np.ascontiguousarray(conf.numpy(), dtype=np.float32)
        if cfg["if_occurance_threshold"]:
            mode, frac = 1, float(cfg["occurance_threshold"])
        elif cfg["if_detected_ratio_threshold"]:
            mode, frac = 0, float(cfg["detected_ratio_threshold"])
        else:
            mode, frac = 2, 0.0
        self.mode = mode
        cap = n // max(int(cfg["min_aggragated_masks"]), 1) + 1
        self.agg = torch.empty((cap, rows.shape[1]), dtype=torch.int32, device=rows.device)
        self.n_groups = np.zeros(1, np.int32)
        self.first, self.size = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
        self.gconf = np.zeros(cap, np.float32)
        self.before, self.after = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
        self.members = np.zeros(n, np.int32)
        self.fval = np.zeros(1, np.float32)
        mc = mc.reshape(-1)
        self.keep = (rows, conf32, label_id, mc, vc)         # referenced by raw pointers below
        d = self.desc = _cabi.Stage2Desc()
        d.rows, d.n_rows, d.n_points = _cabi.ptr(rows), n, int(n_points)
        d.conf, d.label_id, d.mc, d.vc = _cabi.ptr(conf32), _cabi.ptr(label_id), _cabi.ptr(mc), _cabi.ptr(vc if mode == 0 else None)
        d.iou_thres = float(np.float32(cfg["iou_thres"]))
        d.min_aggregated, d.filter_mode, d.filter_fraction = int(cfg["min_aggragated_masks"]), mode, frac
        d.agg_rows, d.cap_groups = _cabi.ptr(self.agg), cap
        d.n_groups, d.group_first, d.group_size, d.group_conf = (_cabi.ptr(self.n_groups), _cabi.ptr(self.first), _cabi.ptr(self.size),
                                                                 _cabi.ptr(self.gconf))
        d.before, d.after, d.members, d.filter_value = _cabi.ptr(self.before), _cabi.ptr(self.after), _cabi.ptr(self.members), _cabi.ptr(self.fval)
        if pairs is not None:       # (inter (n,n) int32, pop (n,) int32) CUDA: the all-pairs step was done elsewhere (dist.sharded_pairs)
            inter_pre, pop_pre = pairs
            if tuple(inter_pre.shape) != (n, n) or tuple(pop_pre.shape) != (n,) or inter_pre.dtype != torch.int32 or pop_pre.dtype != torch.int32:
                raise ValueError("pairs must be ((n,n) int32, (n,) int32)")
            self.pairs = (inter_pre.contiguous(), pop_pre.contiguous())
            d.inter_pre, d.pop_pre = _cabi.ptr(self.pairs[0]), _cabi.ptr(self.pairs[1])

    def result(self):
        from . import _cabi
        g = int(self.n_groups[0])
        _cabi.launch_count += 5 + (12 if g else 0)      # kernels inside the call (popcount x2, intersect, adjacency, ...)
        if g == 0:
            log("masks_to_be_merged", [])
            return None
        off = np.concatenate([[0], np.cumsum(self.size[:g])])
        groups = [self.members[off[k]:off[k + 1]].tolist() for k in range(g)]
        log("masks_to_be_merged", groups)
        if self.mode != 2 and _log.VERBOSE:
            log("detected_ratio_thres_value" if self.mode == 0 else "occurance_thres_value", torch.tensor(self.fval[0]))
        return (self.agg[:g], torch.from_numpy(self.gconf[:g].copy()), [self.labels[int(i)] for i in self.first[:g]], groups,
                torch.from_numpy(self.before[:g].astype(np.int64)), torch.from_numpy(self.after[:g].astype(np.int64)))


def aggregate_filter_fused(rows: torch.Tensor, conf: torch.Tensor, labels: Sequence[str], mc: torch.Tensor,
                           vc: Optional[torch.Tensor], n_points: int, cfg, pairs=None):
    """aggregate_rows + keep_points + the overlap/AND/popcount half of filter_and_select in ONE call
    (bff_stage2_aggregate): the component search, threshold pick and overlap schedule run in C++ next to the
    kernel launches. Returns None when nothing aggregates, else
    (agg_rows CUDA (G,Wd), conf_agg CPU (G,), labels_agg, groups, before (G,) int64, after (G,) int64)."""
    from . import _cabi
    job = _FusedJob(rows, conf, labels, mc, vc, n_points, cfg, pairs)
    _cabi.check(_cabi.lib().bff_stage2_aggregate(_cabi.ctx(rows.device.index or 0), job.desc, _cabi.current_stream()), "bff_stage2_aggregate")
    return job.result()


def aggregate_filter_fused_batch(jobs, cfg):
    """aggregate_filter_fused for several queries of one scene in ONE call (bff_stage2_aggregate_batch): the phases of all
    jobs share their stream synchronisations. jobs: [(rows, conf, labels, mc, vc, n_points), ...] on one device.
    Returns one aggregate_filter_fused result per job (log lines come out job by job, in order)."""
    from . import _cabi
    if not jobs:
        return []
    prepared = [_FusedJob(*j, cfg) for j in jobs]
    descs = (_cabi.Stage2Desc * len(prepared))(*[p.desc for p in prepared])       # copies the structs; the buffers stay owned by `prepared`
    dev = jobs[0][0].device
    _cabi.check(_cabi.lib().bff_stage2_aggregate_batch(_cabi.ctx(dev.index or 0), descs, len(prepared), _cabi.current_stream()),
                "bff_stage2_aggregate_batch")
    return [p.result() for p in prepared]


def select_rows(agg: torch.Tensor, conf_agg: torch.Tensor, labels_agg: Sequence[str], before64: torch.Tensor,
                after64: torch.Tensor, n_points: int, cfg, to_host: bool = True):
    """Row filter + the saved dict (projection_2d_to_3d.py:601-634), torch semantics for the comparisons."""
    sel = (after64 > cfg["remove_small_masks"]) & (after64 > cfg["remove_filtered_masks"] * before64)
    idx = torch.nonzero(sel).reshape(-1)
    # every row kept (the usual case): the buffer itself, no index upload and no gather
    out_rows = agg if idx.numel() == agg.shape[0] else agg[ops.small_to_device(idx, agg.device)]
    ins = ops.unpack_bits(out_rows, n_points)
    log("after filtering", ins.shape)
    return {"ins": ops.to_host(ins) if to_host else ins, "conf": conf_agg[sel], "final_class": [labels_agg[i] for i in idx.tolist()]}, out_rows
