"""Device-resident aggregation / filtering / refinement (bff_stage2_device, bff_refine_device).

The same decisions as aggregate.py / refine.py (tools/projection_2d_to_3d.py:100-301,512-623 and
tools/refinement.py:208-412), but every one of them is taken by a small kernel on the GPU: a call only ENQUEUES work on
the current stream. Counts that are born on the device (aggregated groups, selected rows, matches, final masks) stay
there in a result block; `header()` / `final()` copy the block to the host and synchronise ONCE, when the caller wants
the numbers. A scene therefore runs lift -> aggregate -> filter -> refine without a single host round trip in between.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch

from . import _cabi
from ._cabi import check, current_stream, lib, ptr


def _filter_mode(cfg):
    if cfg["if_occurance_threshold"]:
        return 1, float(cfg["occurance_threshold"])
    if cfg["if_detected_ratio_threshold"]:
        return 0, float(cfg["detected_ratio_threshold"])
    return 2, 0.0


def _scalar_kind(v):
    """torch's promotion for `int64 tensor > python scalar` (projection_2d_to_3d.py:601-623): a Python int compares as
    int64, a Python float in float32. Returns (is_float, int value, float32 value)."""
    if isinstance(v, bool) or isinstance(v, (int, np.integer)):
        return 0, int(v), 0.0
    return 1, 0, float(np.float32(v))


class Stage2Dev:
    """Outputs of one bff_stage2_device call (all on the device until `header()`)."""

    def __init__(self, n_rows, cap, n_points, out_rows, result, labels):
        self.n_rows, self.cap, self.n_points = n_rows, cap, n_points
        self.out_rows, self.result, self.labels = out_rows, result, labels
        self._host = None

    # device-side views the refinement call consumes without knowing the counts
    @property
    def n_selected_dev(self) -> torch.Tensor:
        return self.result[_cabi.S2R_N_SELECTED:_cabi.S2R_N_SELECTED + 1]

    @property
    def selected_conf_dev(self) -> torch.Tensor:
        o = _cabi.S2R_HEADER + 6 * self.cap
        return self.result[o:o + self.cap].view(torch.float32)

    def header(self) -> dict:
        """Result block on the host (one synchronisation, cached)."""
        if self._host is None:
            r = self.result.cpu().numpy()
            c, o = self.cap, _cabi.S2R_HEADER
            g, m = int(r[_cabi.S2R_N_GROUPS]), int(r[_cabi.S2R_N_SELECTED])
            arr = lambda k, n, dt=np.int32: r[o + k * c:o + k * c + n].view(dt).copy()
            seg = r[o + 7 * c:o + 7 * c + g + 1]
            members = r[o + 8 * c + 1:o + 8 * c + 1 + self.n_rows]
            err = int(r[_cabi.S2R_ERROR])
            if err & 1:
                raise IndexError("the filter fraction selects an index outside the unique vote values "
                                 "(the reference raises IndexError at projection_2d_to_3d.py:516/576)")
            if err & 2:
                raise _cabi.BffError("bff_stage2_device: a scratch bound (mc_cap / vc_cap / cap_groups) was exceeded")
            self._host = dict(
                n_groups=g, n_selected=m, n_overlap=int(r[_cabi.S2R_N_OVERLAP]), n_unique=int(r[_cabi.S2R_N_UNIQUE]),
                filter_value=float(r[_cabi.S2R_FILTER_VALUE:_cabi.S2R_FILTER_VALUE + 1].view(np.float32)[0]),
                group_first=arr(0, g), group_size=arr(1, g), group_conf=arr(2, g, np.float32), before=arr(3, g), after=arr(4, g),
                selected=arr(5, m), selected_conf=arr(6, m, np.float32),
                groups=[members[int(seg[k]):int(seg[k + 1])].tolist() for k in range(g)])
        return self._host

    def saved_dict(self, to_host: bool = True, keep_rows: bool = False):
        """The dict projection_2d_to_3d.py:630-634 saves (or its empty form), plus `reached_filter`."""
        from . import ops, scene_io
        from ._log import log
        h = self.header()
        log("masks_to_be_merged", h["groups"])
        if h["n_unique"]:
            log("filter_thres_value", torch.tensor(h["filter_value"]))
        if h["n_groups"] == 0:
            log("No 3d masks detected after aggregation")
            return scene_io.empty_stage2(), False
        log("before filtering", torch.Size([h["n_groups"], self.n_points]))
        m = h["n_selected"]
        rows = self.out_rows[:m]
        ins = ops.unpack_bits(rows, self.n_points)
        log("after filtering", ins.shape)
        out = {"ins": ops.to_host(ins) if to_host else ins, "conf": torch.from_numpy(h["selected_conf"].copy()),
               "final_class": [self.labels[int(h["group_first"][int(k)])] for k in h["selected"]]}
        if keep_rows:
            out["rows"] = rows
        return out, True


def stage2_device(rows: torch.Tensor, conf: torch.Tensor, labels: Sequence[str], mc: torch.Tensor, vc: Optional[torch.Tensor],
                  n_points: int, cfg, mc_cap: Optional[int] = None, vc_cap: int = 0, label_id: Optional[torch.Tensor] = None,
                  vote_point0: int = 0, n_vote_points: Optional[int] = None, comm=None, lane: int = 0) -> Stage2Dev:
    """aggregate + point filter + solve_overlapping + row filter of one query's lifted rows, enqueued without any
    synchronisation. rows (n,Wd) int32 CUDA; conf (n,) float32 CUDA; mc / vc int32 CUDA vote vectors (over ALL points
    when the rows are one rank's point slice: then `vote_point0` is the slice's first point and `n_points` its length);
    mc_cap / vc_cap bound the counts (masks of the query / viewed frames)."""
    n = rows.shape[0]
    if n == 0:
        raise ValueError("stage2_device needs at least one lifted row (the caller handles 'nothing lifted')")
    if conf.dtype != torch.float32 or not conf.is_cuda:
        raise TypeError("conf must be a float32 CUDA tensor")
    min_agg = int(cfg["min_aggragated_masks"])
    if min_agg < 1:
        raise ValueError("stage2_device needs min_aggragated_masks >= 1")
    if label_id is None and not (isinstance(labels, (list, tuple)) and labels.count(labels[0]) == n):
        ids: dict = {}
        label_id = torch.tensor([ids.setdefault(l, len(ids)) for l in labels], dtype=torch.int32).to(rows.device)
    mode, frac = _filter_mode(cfg)
    cap = n // min_agg + 1
    dev = rows.device
    out_rows = torch.empty((cap, rows.shape[1]), dtype=torch.int32, device=dev)
    result = torch.empty((_cabi.stage2_dev_result_words(n, cap),), dtype=torch.int32, device=dev)
    mc = mc.reshape(-1)
    d = _cabi.Stage2DevDesc()
    d.rows, d.n_rows, d.n_points = ptr(rows), n, int(n_points)
    d.conf, d.label_id, d.mc, d.vc = ptr(conf), ptr(label_id), ptr(mc), ptr(vc if mode == 0 else None)
    d.n_vote_points, d.vote_point0 = int(mc.numel() if n_vote_points is None else n_vote_points), int(vote_point0)
    d.iou_thres, d.min_aggregated, d.filter_mode, d.filter_fraction = float(np.float32(cfg["iou_thres"])), min_agg, mode, frac
    d.small_is_float, d.small_i, d.small_f = _scalar_kind(cfg["remove_small_masks"])
    d.ratio_is_float, d.ratio_i, d.ratio_f = _scalar_kind(cfg["remove_filtered_masks"])
    d.mc_cap, d.vc_cap = int(n if mc_cap is None else mc_cap), int(vc_cap)
    d.out_rows, d.cap_groups, d.result, d.comm = ptr(out_rows), cap, ptr(result), comm
    check(lib().bff_stage2_device(_cabi.ctx(dev.index or 0, lane), d, current_stream()), "bff_stage2_device")
    _cabi.launch_count += 17 if mode != 2 else 13
    return Stage2Dev(n, cap, int(n_points), out_rows, result, labels)


class RefineDev:
    """Outputs of one bff_refine_device call."""

    def __init__(self, desc, keep, n1, cap2, n_points, result, final_rows, s2_out, query):
        self.desc, self._keep = desc, keep
        self.n1, self.cap2, self.n_points = n1, cap2, n_points
        self.result, self.final_rows, self.s2_out, self.query = result, final_rows, s2_out, query
        self._host = None
        self._spec = None

    def _off(self):
        c, n1 = self.cap2, self.n1
        o = {}
        p = _cabi.RFR_HEADER
        for name, n in (("sims", 2 * c), ("first_match", c), ("best", c), ("match", c), ("match_iou", c), ("conf_out", c),
                        ("idx2", c), ("seg2", c + 1), ("other", n1), ("final_src", n1 + c), ("final_conf", n1 + c)):
            o[name] = (p, n)
            p += n
        return o

    @property
    def n_final_dev(self) -> torch.Tensor:
        return self.result[_cabi.RFR_N_FINAL:_cabi.RFR_N_FINAL + 1]

    def enqueue_unpack(self, cap_rows: int = 64):
        """Speculatively expand the first `cap_rows` final rows to bool on the device (bff_unpack_bits_dc) before the
        count is known on the host; `final()` falls back to an exact unpack when there are more."""
        cap_rows = min(cap_rows, self.n1 + self.cap2)
        buf = torch.empty((cap_rows, self.n_points), dtype=torch.uint8, device=self.result.device)
        check(lib().bff_unpack_bits_dc(ptr(self.final_rows), ptr(self.n_final_dev), cap_rows, self.n_points, ptr(buf), current_stream()),
              "bff_unpack_bits_dc")
        _cabi.launch_count += 1
        self._spec = buf

    def header(self) -> dict:
        if self._host is None:
            r = self.result.cpu().numpy()
            o = self._off()
            m, k = int(r[_cabi.RFR_N2_OUT]), int(r[_cabi.RFR_N_FINAL])
            if int(r[_cabi.RFR_ERROR]) & 1:
                raise IndexError("the similarity percentile selects an index outside the unique similarities "
                                 "(the reference raises IndexError at refinement.py:324)")
            get = lambda name, n, dt=np.int32: r[o[name][0]:o[name][0] + n * (2 if dt == np.float64 else 1)].view(dt).copy()
            self._host = dict(
                n2_in=int(r[_cabi.RFR_N2_IN]), n2_out=m, n_or=int(r[_cabi.RFR_N_OR]), n_other=int(r[_cabi.RFR_N_OTHER]), n_final=k,
                sim_thres=float(r[_cabi.RFR_SIM_THRES:_cabi.RFR_SIM_THRES + 2].view(np.float64)[0]),
                sims=get("sims", m, np.float64), first_match=get("first_match", int(r[_cabi.RFR_N2_IN])),
                best=get("best", int(r[_cabi.RFR_N2_IN])), match=get("match", m), match_iou=get("match_iou", m, np.float32),
                conf_out=get("conf_out", m, np.float32), other=get("other", int(r[_cabi.RFR_N_OTHER])),
                final_src=get("final_src", k), final_conf=get("final_conf", k, np.float32))
        return self._host

    def set_sims(self, sims: Sequence[float]):
        """Replace the similarities of the matched pairs (one per remaining stage-2 mask, refinement.py:297-302) before
        `select`: callers whose similarity function is expensive (CLIP) evaluate it for the MATCHED labels only, like the
        reference, instead of filling the whole per-mask table up front."""
        v = np.asarray(list(sims), dtype=np.float64)
        if v.size:
            o = self._off()["sims"][0]
            self.result[o:o + 2 * v.size].copy_(torch.from_numpy(v.view(np.int32)).to(self.result.device))
        self._host = None

    def select(self, sim_thres: float):
        """Pass 2 with the class-global threshold of refinement.py:317-324 (after bff_refine_device in select_mode 0)."""
        dev = self.result.device
        if self.final_rows is None:
            self.final_rows = torch.empty((self.n1 + self.cap2, self.s2_out.shape[1]), dtype=torch.int32, device=dev)
            self.desc.final_rows = ptr(self.final_rows)
        check(lib().bff_refine_select_device(_cabi.ctx(dev.index or 0), self.desc, float(sim_thres), current_stream()), "bff_refine_select_device")
        _cabi.launch_count += 2
        self._host = None

    def final(self, to_host: bool = True) -> dict:
        """The dict refinement.py:422-426 saves: {"ins": bool (k,N), "conf": float32 (k,), "final_class": [query] * k},
        list forms when nothing is left (:347-358,:402-409)."""
        from . import ops
        h = self.header()
        k = h["n_final"]
        if k == 0:
            return {"ins": [], "conf": [], "final_class": []}
        if self._spec is not None and k <= self._spec.shape[0]:
            ins = self._spec[:k].view(torch.bool)
        else:
            ins = ops.unpack_bits(self.final_rows[:k], self.n_points)
        return {"ins": ops.to_host(ins) if to_host else ins, "conf": torch.from_numpy(h["final_conf"].copy()),
                "final_class": [self.query] * k}


def refine_device(s1_rows: torch.Tensor, s2_rows: torch.Tensor, n2, s2_conf: torch.Tensor, is_query: torch.Tensor, sim: torch.Tensor,
                  n_points: int, cfg, query: str, select_mode: int = 1, sim_thres: float = 0.0, comm=None, lane: int = 0) -> RefineDev:
    """Pass 1 (+ pass 2 unless select_mode 0) of tools/refinement.py for one scene, enqueued without synchronisation.
    s1_rows (n1,Wd) int32 CUDA, modified in place (:248); s2_rows (cap2,Wd) of which the first n2 count -- `n2` is an int
    or a 1-element int32 CUDA tensor (Stage2Dev.n_selected_dev); s2_conf (cap2,) float32 CUDA; is_query (n1,) uint8 CUDA:
    stage-1 label == query with underscores; sim (n1,) float64 CUDA: text similarity of the query and each stage-1 label."""
    n1, cap2 = s1_rows.shape[0], s2_rows.shape[0]
    if n1 == 0 or cap2 == 0:
        raise ValueError("refine_device needs stage-1 masks and room for stage-2 masks")
    dev = s1_rows.device
    if s2_conf.dtype != torch.float32 or is_query.dtype != torch.uint8 or sim.dtype != torch.float64:
        raise TypeError("s2_conf float32, is_query uint8, sim float64 expected")
    result = torch.empty((_cabi.refine_dev_result_words(n1, cap2),), dtype=torch.int32, device=dev)
    s2_out = torch.empty_like(s2_rows)
    final_rows = torch.empty((n1 + cap2, s1_rows.shape[1]), dtype=torch.int32, device=dev) if select_mode != 0 else None
    d = _cabi.RefineDevDesc()
    d.s1_rows, d.n1, d.s2_rows, d.cap2, d.n_points = ptr(s1_rows), n1, ptr(s2_rows), cap2, int(n_points)
    if isinstance(n2, torch.Tensor):
        d.n2_dev, d.n2_host = ptr(n2), 0
    else:
        d.n2_dev, d.n2_host = None, int(n2)
    d.s2_conf, d.stage1_iou_thres = ptr(s2_conf), float(np.float32(cfg["stage1_iou_thres"]))
    d.is_query, d.sim = ptr(is_query), ptr(sim)
    d.select_mode, d.sim_percentile, d.sim_thres = int(select_mode), float(cfg["refinment_sim_percentile"]), float(sim_thres)
    d.refine_iou_thres = float(np.float32(cfg["refiment_iou_thres"]))
    d.s2_out, d.final_rows, d.result, d.comm = ptr(s2_out), ptr(final_rows), ptr(result), comm
    check(lib().bff_refine_device(_cabi.ctx(dev.index or 0, lane), d, current_stream()), "bff_refine_device")
    _cabi.launch_count += 11 + (2 if select_mode else 0) + (0 if isinstance(n2, torch.Tensor) else 1)
    return RefineDev(d, (s1_rows, s2_rows, n2, s2_conf, is_query, sim), n1, cap2, int(n_points), result, final_rows, s2_out, query)


def stage1_tables(stage1_labels: Sequence[str], query: str, sim_fn, device):
    """The two per-mask tables of the stage-1 side that pass 1/2 look up on the device: is_query (label == query with
    underscores, refinement.py:293) and sim (sim_fn(query, label), :297-302; one evaluation per DISTINCT label)."""
    q_ = query.replace(" ", "_")
    cache: dict = {}
    sims = np.asarray([cache[l] if l in cache else cache.setdefault(l, float(sim_fn(query, l))) for l in stage1_labels], dtype=np.float64)
    isq = np.asarray([l == q_ for l in stage1_labels], dtype=np.uint8)
    return torch.from_numpy(isq).to(device), torch.from_numpy(sims).to(device)
