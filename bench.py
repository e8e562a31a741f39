#!/usr/bin/env python
"""bench.py -- throughput of the BeyondFF lifting + refinement hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (our arm; torchrun launches N>1)
    python bench.py --impl reference --gpus N --steps K ...   (the reference's CPU path, host cores)

A "step" is one pass of the whole hot path over one synthetic ScanNet200-shaped scene
(BASELINE.json configs[1]: 250k points, 300 frames 640x480, 1 query, 200 stage-1 masks):
RLE 2D masks -> packed bits, projection + visibility + mask gather + votes for every frame
(bff_lift_scene), all-pairs IoU + aggregation, point filter, overlap resolution, then the
stage-1/stage-2 refinement of that scene. `value` = point-view projections (N x F, each
(frame, point) counted once) per second of the WHOLE step with inputs resident in HBM; `e2e`
is the same through the host API with pinned host buffers in and the final masks out.
The whole step is device-resident (no host round trip between the stages) and `value` replays it as ONE CUDA graph.
With N > 1 the headline line is ONE scene (BASELINE.json configs[3] at 1000 frames, "c4s") strong-scaled over the N GPUs:
frames interleaved over the ranks, the lift kernel stores each peer's point slice of the rows straight into that peer's
buffer (CUDA IPC over NVLink), votes are all-reduced, aggregation / filter / refinement run point-sharded with int32
all-reduces inside, and the final rows are all-gathered; rank 0 re-runs the same scene on one GPU in the same run and
asserts bit-identical final masks (`single_gpu_ms`). `replicas` (one c2 scene per GPU, no collective) and
`frame_sharded_c2` are secondaries of the same line.

Only the cpu_baseline / --impl reference legs touch oracle/ (the reference's numpy path
restated, oracle/reference_port.py); the GPU arm raises if libbff.so is missing.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))

import numpy as np
import torch

WORKLOADS = {
    # BASELINE.json configs: N points, frames PER RANK, mask size W x H, stage-1 masks, boxes, text queries,
    # native depth size (0 = same as the masks; else the lift kernel resamples like cv2.resize, :436)
    "c1": dict(n_points=150_000, n_frames=100, width=640, height=480, n_stage1=50, n_boxes=14),
    "c2": dict(n_points=250_000, n_frames=300, width=640, height=480, n_stage1=200, n_boxes=14),
    # configs[2]: the c2 scene with 20 concurrent text queries (shared projection, per-query votes and rows)
    "c3": dict(n_points=250_000, n_frames=300, width=640, height=480, n_stage1=200, n_boxes=40, n_queries=20),
    # configs[3]: 2M points, 2000 frames over 8 ranks (250 per rank), 1296x968 masks over 640x480 depth, 500 stage-1 masks
    "c4": dict(n_points=2_000_000, n_frames=250, width=1296, height=968, depth_width=640, depth_height=480,
               n_stage1=500, n_boxes=14, room=(20.0, 15.0, 3.0)),
    # configs[3] as ONE scene (strong scaling over the GPUs): all 2000 frames
    "c4s": dict(n_points=2_000_000, n_frames=2000, width=1296, height=968, depth_width=640, depth_height=480,
                n_stage1=500, n_boxes=14, room=(20.0, 15.0, 3.0)),
    # configs[2] as ONE scene (20 queries over the c2 scene; = c3, named for symmetry)
    "tiny": dict(n_points=20_000, n_frames=12, width=320, height=240, n_stage1=20, n_boxes=9),
    "tiny3": dict(n_points=20_000, n_frames=12, width=320, height=240, depth_width=160, depth_height=120,
                  n_stage1=20, n_boxes=9, n_queries=3),
}
QUERY = "office chair"


def query_names(n_queries: int):
    """QUERY first, then other ScanNet200 class names (spaces instead of underscores, like --cls)."""
    from bff_b200.refine import STAGE1_CLASS_NAMES
    names = [QUERY] + [c.replace("_", " ") for c in STAGE1_CLASS_NAMES if c.replace("_", " ") != QUERY]
    return names[:n_queries]


# ----------------------------------------------------------------------------- synthetic inputs
def build_inputs(workload: str, rank: int, world: int, device, max_frames=None, seed: int = 2000, strong: bool = False):
    """In-memory inputs of one scene in the reference's wire formats (RLE 2D masks per frame and query,
    uint16 depth at its native size, float64 poses/intrinsics, RLE stage-1 masks). Frames [rank*F, (rank+1)*F) of a
    (F*world)-frame trajectory are rendered on `device` (input fabrication, untimed)."""
    from bff_b200 import synth
    from bff_b200.refine import STAGE1_CLASS_NAMES
    from bff_b200.rle import rle_encode_rows
    w = WORKLOADS[workload]
    f_rank = w["n_frames"]
    fsel = None
    if strong:
        # ONE scene of w["n_frames"] frames (strong scaling): rank r takes frames r, r + world, ... -- interleaved, so that
        # every rank sees the whole trajectory and the ranks' lifting times are equal (contiguous blocks differed by 25 %:
        # some stretches of the trajectory look at much more of the cloud than others)
        fsel = list(range(rank, w["n_frames"], world))
        f_rank = len(fsel)
    nq = w.get("n_queries", 1)
    queries = query_names(nq)
    n_render = f_rank if max_frames is None else min(max_frames, f_rank)   # the CPU arms only need their sample
    spec = synth.SceneSpec(n_points=w["n_points"], n_frames=w["n_frames"] if strong else f_rank * world, width=w["width"], height=w["height"],
                           depth_width=w.get("depth_width", 0), depth_height=w.get("depth_height", 0),
                           n_boxes=w["n_boxes"], n_stage1=w["n_stage1"], seed=seed, n_queries=nq,
                           room=w.get("room", (6.0, 5.0, 2.8)))
    sc = synth.make_scene(spec)
    if fsel is None:
        fsel = list(range(rank * f_rank, (rank + 1) * f_rank))
    hgt, wid = spec.height, spec.width
    dh, dw = (spec.depth_height or hgt), (spec.depth_width or wid)
    depth = np.zeros((f_rank, dh, dw), np.uint16)
    records = [[] for _ in queries]
    for k, f in enumerate(fsel[:n_render]):
        mm, obj = synth.render_frame(sc, f, device)
        depth[k] = mm.cpu().numpy().astype(np.uint16)
        rng = np.random.default_rng(spec.seed * 7919 + f)
        obj = obj.cpu()
        for qi, q in enumerate(queries):
            ms = synth.frame_masks(sc, f, obj, qi, rng)
            if ms:
                records[qi].append({"frame_id": f"{sc.frame_ids[f]}.jpg",
                                    "segmented_frame_masks": rle_encode_rows(torch.stack(ms).reshape(len(ms), -1)),
                                    "confidences": torch.as_tensor(rng.uniform(0.2, 0.45, len(ms)).astype(np.float32)),
                                    "labels": [q] * len(ms)})
    rng = np.random.default_rng(spec.seed + 7)
    m1, c1, conf1 = synth.stage1_masks(sc, queries, STAGE1_CLASS_NAMES, rng)
    stage1 = {"ins": rle_encode_rows(torch.from_numpy(m1)), "conf": torch.from_numpy(conf1),
              "final_class": torch.from_numpy(c1)}
    poses = np.stack([sc.poses[f] for f in fsel])
    pinv = np.stack([np.linalg.inv(p) for p in poses])
    return dict(spec=spec, xyz=sc.xyz, K=sc.K, poses=poses, pinv=pinv, depth=depth, records=records[0],
                records_by_query=records, queries=queries,
                frame_ids=[str(sc.frame_ids[f]) for f in fsel], stage1=stage1, cfg=dict(synth.DEFAULT_CFG))


def fake_similarity(query: str, label: str) -> float:
    """CLIP text similarity is an INPUT of the path (refinement.py:93-117); a deterministic table stands in."""
    import hashlib
    h = int.from_bytes(hashlib.sha1((query + "|" + label).encode()).digest()[:4], "little")
    return 0.5 + (h % 1000) / 2000.0


# ----------------------------------------------------------------------------- our arm
class GpuScene:
    """Host-staged (pinned) and device-resident forms of one rank's inputs + the step itself.
    Q text queries share one lifting launch (projection, visibility, viewed counts); rows and votes are per query."""

    def __init__(self, inp, device, dist_ctx=None):
        from bff_b200 import lift as lift_mod, scene_io
        from bff_b200.rle import pairs_for_device
        self.inp, self.dev, self.dist = inp, device, dist_ctx
        spec = inp["spec"]
        self.n, self.h, self.w = spec.n_points, spec.height, spec.width
        self.f = inp["depth"].shape[0]
        self.native = tuple(inp["depth"].shape[1:])                 # depth PNG size; != (h, w): resampled in the lift kernel
        self.names = list(inp["queries"])
        self.nq = len(self.names)
        self.queries = [lift_mod.QueryMasks.from_file_records(nm, recs) for nm, recs in zip(self.names, inp["records_by_query"])]
        self.q = self.queries[0]
        self.frame_ids, self.viewed, self.per_frame = lift_mod.plan_frames(self.queries, inp["frame_ids"])
        order = {fid: i for i, fid in enumerate(inp["frame_ids"])}
        perm = [order[fid] for fid in self.frame_ids]           # mask frames first, then view-only frames
        rles, mask_query, offsets = lift_mod.mask_layout(self.per_frame, self.queries, 0, len(self.frame_ids))
        counts, pair_off, length = pairs_for_device(rles)
        assert length == self.h * self.w or not rles
        self.n_masks = len(rles)
        self.row_index = lift_mod.query_row_index(self.per_frame, self.queries)   # per query: its rows in file order
        self.conf_q, self.labels_q = [], []
        for q in self.queries:
            c, l = q.flat_conf_and_labels()
            self.conf_q.append(torch.tensor(c) if c else [])    # :484-486, parsed once with the file
            self.labels_q.append(l)
        self.conf_list, self.labels = self.conf_q[0], self.labels_q[0]
        self.my_queries = list(range(self.nq))                  # frame-sharded runs split the queries over the ranks
        s1_counts, s1_off, s1_len = pairs_for_device(list(inp["stage1"]["ins"]))
        assert s1_len == self.n

        def pin(a):
            t = torch.from_numpy(np.ascontiguousarray(a))
            return t.pin_memory() if torch.cuda.is_available() else t
        x32 = inp["xyz"].astype(np.float32)
        assert np.array_equal(x32.astype(np.float64), inp["xyz"])
        x4 = np.zeros((self.n, 4), np.float32)
        x4[:, :3] = x32
        self.plan = None
        if dist_ctx is not None and dist_ctx.world_size > 1:
            from bff_b200.dist import ShardPlan
            self.plan = ShardPlan(self.n, dist_ctx.world_size, dist_ctx.rank)
            x4 = self.plan.padded_cloud(x4)        # NaN points behind the cloud: every point slice has the same width
        # host side (pinned): exactly what bff's loaders hand to the device
        self.host = dict(xyz=pin(x4), pinv=pin(inp["pinv"][perm]), depth=pin(inp["depth"][perm].view(np.int16)),
                         counts=pin(counts), pair_off=pin(pair_off), offsets=pin(np.asarray(offsets, np.int32)),
                         viewed=pin(np.asarray(self.viewed, np.uint8)), s1_counts=pin(s1_counts), s1_pair_off=pin(s1_off))
        if self.nq > 1:
            self.host["mask_query"] = pin(np.asarray(mask_query, np.int32))
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host.values())
        # device-resident decisions (bff_b200/devpipe.py): per query the lifted rows' confidences and the two stage-1 tables
        # (is the label the query's class; CLIP-style similarity of query and label -- an INPUT of the path) travel with
        # the scene's other inputs
        from bff_b200 import devpipe
        from bff_b200.refine import STAGE1_CLASS_NAMES
        self.device_decisions = os.environ.get("BFF_DEVICE_DECISIONS", "1") != "0"
        self.use_graph = os.environ.get("BFF_STEP_GRAPH", "1") != "0"        # replay the step as one CUDA graph
        s1_labels = [STAGE1_CLASS_NAMES[int(i)] for i in inp["stage1"]["final_class"]]
        for qi, name in enumerate(self.names):
            q_ = name.replace(" ", "_")
            sims = {}
            self.host[f"conf{qi}"] = pin(self.conf_q[qi].numpy() if len(self.conf_q[qi]) else np.zeros(1, np.float32))
            self.host[f"isq{qi}"] = pin(np.asarray([l == q_ for l in s1_labels], np.uint8))
            self.host[f"sim{qi}"] = pin(np.asarray([sims[l] if l in sims else sims.setdefault(l, fake_similarity(name, l)) for l in s1_labels], np.float64))
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in self.host.values())
        self.row_index_dev = None
        if self.nq > 1 and torch.cuda.is_available():
            self.row_index_dev = [torch.tensor(ix, dtype=torch.long, device=device) for ix in self.row_index]
            self.row_index_dev32 = [t.to(torch.int32) for t in self.row_index_dev]
        self.lift_events = None
        if torch.cuda.is_available():
            # recorded inside bff_lift_scene around its memsets + kernel: the launch itself, without host gaps
            self.lift_events = (_event(), _event())

    def masks2d(self, d):
        """The 2D masks of all frames as packed bit rows, decoded on the device from the RLE wire format."""
        from bff_b200 import ops
        return ops.rle_pairs_to_bits_dev(d["counts"], d["pair_off"], self.h * self.w) if self.n_masks else None

    def stage1_rows(self, d):
        from bff_b200 import ops
        return ops.rle_pairs_to_bits_dev(d["s1_counts"], d["s1_pair_off"], self.n)

    def upload(self):
        return {k: v.to(self.dev, non_blocking=True) for k, v in self.host.items()}

    def lift(self, d, masks2d, timing=False, frames_per_chunk=0):
        from bff_b200 import ops
        return ops.lift_scene(d["xyz"], d["pinv"], self.inp["K"], d["depth"], self.h, self.w, masks2d=masks2d,
                              mask_offset=d["offsets"] if self.n_masks else None,
                              mask_query=d.get("mask_query") if self.n_masks else None, n_queries=self.nq,
                              frame_viewed=d["viewed"], depth_thresh=0.08, depth_scale=1000.0, depth_range=(0.0, 65.535),
                              native_size=self.native, frames_per_chunk=frames_per_chunk,
                              timing_events=self.lift_events if timing else None)

    def step(self, d, timers=None, to_host=True):
        """One pass of the hot path on device-resident inputs `d`. Returns the final dict of query 0 (all queries'
        results are in self.last_finals); its (k,N) bool masks are copied to the host (pinned) when `to_host`,
        else they stay in HBM like the inputs."""
        from bff_b200 import ops, refine, stage2
        inp, cfg = self.inp, self.inp["cfg"]
        ev = (lambda: None) if timers is None else (lambda: timers.append(_event()))
        self._timing = timers is not None
        if self.dist is not None and self.dist.world_size > 1 and getattr(self, "comm", None) is not None:
            return self.step_sharded(d, ev, to_host)
        if self.device_decisions and not (self.dist is not None and self.dist.world_size > 1) and cfg["min_aggragated_masks"] >= 1:
            return self.step_device(d, ev, to_host)
        ev()
        masks2d = self.masks2d(d)
        ev()
        out = self.lift(d, masks2d, timing=timers is not None)
        ev()
        rows, mc, vc = out.rows, out.masked_counts, out.viewed_counts
        sharded = self.dist is not None and self.dist.world_size > 1
        from bff_b200 import dist as dist_mod
        if sharded:
            rows, mc, vc = dist_mod.exchange(self.dist, rows, self.rows_per_rank, mc, vc)
        ev()
        s1 = self.stage1_rows(d)

        def job(qi):
            if sharded:
                conf_list, labels, idx = self.all_conf[qi], self.all_labels[qi], self.all_index_dev[qi]
            else:
                conf_list, labels = self.conf_q[qi], self.labels_q[qi]
                idx = self.row_index_dev[qi] if self.row_index_dev is not None else None
            return (rows if idx is None else rows[idx], conf_list, labels, mc[qi], vc, self.n)

        def aggregate(qi):
            rows_q, conf_list, labels = job(qi)[:3]
            pairs = None
            if sharded and self.nq == 1 and rows_q.shape[0] >= dist_mod.PAIRS_MIN_ROWS:
                pairs = dist_mod.sharded_pairs(self.dist, rows_q)    # all-pairs IoU: one row block per rank + all-gather
            res2, _ = stage2.aggregate_and_filter(rows_q, conf_list, labels, mc[qi], vc, self.n, cfg, keep_rows=True, to_host=to_host,
                                                  pairs=pairs)
            return res2

        def refine_query(qi, res2):
            name = self.names[qi]
            # pass 1 ORs matched stage-1 rows in place (refinement.py:248): every query starts from the decoded originals,
            # as every reference process (one per --cls) does
            p1 = refine.pass1_scene(inp["stage1"], res2, name, cfg["stage1_iou_thres"], fake_similarity, self.dev,
                                    s1_rows=s1.clone() if self.nq > 1 else s1)
            thr = refine.similarity_threshold([p1["sims"]], cfg["refinment_sim_percentile"]) if len(p1["sims"]) else 0.0
            return refine.pass2_scene(p1, name, cfg["refiment_iou_thres"], thr, to_host=to_host)

        if True:
            if len(self.my_queries) > 1 and os.environ.get("BFF_AGG_BATCH", "1") != "0":
                # all queries advance phase by phase inside ONE bff_stage2_aggregate_batch call (shared synchronisations)
                per_query = [r for r, _ in stage2.aggregate_and_filter_batch([job(qi) for qi in self.my_queries], cfg, keep_rows=True,
                                                                             to_host=to_host)]
            else:
                per_query = [aggregate(qi) for qi in self.my_queries]
            ev()
            if len(self.my_queries) > 1 and os.environ.get("BFF_AGG_BATCH", "1") != "0":
                # pass 1 of all queries in ONE bff_refine_match_batch call; every query ORs into its own copy of the stage-1 rows
                names = [self.names[qi] for qi in self.my_queries]
                p1s = refine.pass1_scene_batch(inp["stage1"], per_query, names, cfg["stage1_iou_thres"], fake_similarity, self.dev,
                                               [s1.clone() for _ in names])
                finals = []
                for name, p1 in zip(names, p1s):
                    thr = refine.similarity_threshold([p1["sims"]], cfg["refinment_sim_percentile"]) if len(p1["sims"]) else 0.0
                    finals.append(refine.pass2_scene(p1, name, cfg["refiment_iou_thres"], thr, to_host=to_host))
            else:
                finals = [refine_query(qi, r2) for qi, r2 in zip(self.my_queries, per_query)]
            ev()
        self.last_stage2 = per_query[0] if per_query else None
        self.last_finals = finals
        return finals[0] if finals else None


def _step_device(self, d, ev, to_host=True):
    """The same step with every decision taken on the device (bff_b200/scene_step.py: bff_stage2_device /
    bff_refine_device): the whole chain lift -> aggregate -> filter -> refine is enqueued without a host round trip -- or
    replayed as ONE CUDA graph when `use_graph` is set (the graph is captured on the first call with a given set of input
    buffers). The only synchronisation is the one that brings the final counts back, after the last kernel."""
    from bff_b200.scene_step import SceneStep
    if getattr(self, "_scene_step", None) is None:
        self._scene_step = SceneStep(n_points=self.n, height=self.h, width=self.w, intrinsics=self.inp["K"], native_size=self.native,
                                     names=self.names, labels_q=self.labels_q, cfg=self.inp["cfg"], n_stage1=len(self.inp["stage1"]["ins"]),
                                     n_masks=self.n_masks, n_frames=self.f,
                                     row_index=self.row_index_dev32 if self.row_index_dev is not None else None, device=self.dev,
                                     lanes=int(os.environ.get("BFF_QUERY_LANES", "4")))
    st = self._scene_step
    timing = getattr(self, "_timing", False)
    if st.has_graph(d) and not timing:
        finals, state = st.run(d, to_host)
    else:
        st.timing_events = self.lift_events if timing else None
        marks = (lambda: ev()) if timing else None
        finals, state = st.run(d, to_host, marks)
        st.timing_events = None
    s2s = [c[0] for c in state["chains"]]
    self.last_stage2_dev = s2s[0] if s2s else None
    self.last_refine_dev = state["chains"][0][1] if state["chains"] else None
    self.last_finals = finals
    return finals[0] if finals else None


GpuScene.step_device = _step_device


def _graph_inputs(self, d):
    """Record the device-decision step over the buffers of `d` as ONE CUDA graph (later steps on `d` replay it)."""
    if not (self.use_graph and self.device_decisions and self.inp["cfg"]["min_aggragated_masks"] >= 1) or (self.dist is not None and self.dist.world_size > 1):
        return False
    self.step(d, to_host=False)                       # creates the SceneStep, warms workspaces and attributes
    self._scene_step.capture(d)
    return True


GpuScene.graph_inputs = _graph_inputs


def _step_sharded(self, d, ev, to_host=True):
    """One scene on G GPUs: rank r lifts its block of the frames over all points (rows written in G column slices), ONE
    exchange hands every rank all rows restricted to its slice of the points plus the all-reduced votes, then every rank
    runs the device-resident pipelines on its slice; their additive integer results are all-reduced inside the calls
    and all ranks take the same decisions (bff_b200/dist.py). No host round trip before the final counts."""
    from bff_b200 import devpipe, dist as dist_mod, ops
    inp, cfg, plan, comm = self.inp, self.inp["cfg"], self.plan, self.comm
    sw, world = plan.slice_words, plan.world
    ev()
    masks2d = self.masks2d(d)
    ev()
    votes = torch.empty((self.nq + 1, plan.n_padded), dtype=torch.int32, device=self.dev)
    peers = getattr(self, "peers", None)
    if peers is not None:
        # fused exchange: my receive buffer zeroed, everybody told (one-word all-reduce = barrier on the stream), then the
        # lifting kernels store their row slices straight into the owners' buffers over NVLink
        peers.local.zero_()
        comm.allreduce_(peers.flag)
    out = ops.lift_scene(d["xyz"], d["pinv"], inp["K"], d["depth"], self.h, self.w, masks2d=masks2d,
                         mask_offset=d["offsets"] if self.n_masks else None, mask_query=d.get("mask_query") if self.n_masks else None,
                         n_queries=self.nq, frame_viewed=d["viewed"], depth_thresh=0.08, depth_scale=1000.0, depth_range=(0.0, 65.535),
                         native_size=self.native, row_slices=(world, sw), votes=votes,
                         peer_rows=peers.addrs if peers is not None else None, peer_row0=self.row0,
                         timing_events=self.lift_events if getattr(self, "_timing", False) else None)
    ev()
    comm.allreduce_(votes)                           # (with the fused exchange: also the barrier behind the peers' stores)
    rows = peers.local if peers is not None else comm.exchange_rows(out.rows, self.rows_per_rank)
    ev()
    n1 = len(inp["stage1"]["ins"])
    s1 = ops.rows_slice(self.stage1_rows(d), plan.rank * sw, sw)
    s2s = []
    for qi in range(self.nq):
        rows_q = rows if self.all_index_dev32[qi] is None else ops.rows_gather(rows, self.all_index_dev32[qi])
        if rows_q.shape[0] == 0:
            s2s.append(None)
            continue
        s2s.append(devpipe.stage2_device(rows_q, self.all_conf_dev[qi], self.all_labels[qi], votes[qi], votes[self.nq], plan.slice_points, cfg,
                                         mc_cap=rows_q.shape[0], vc_cap=self.f_total, vote_point0=plan.point0, n_vote_points=self.n,
                                         comm=comm.handle))
    ev()
    rfs = []
    for qi, s2 in enumerate(s2s):
        s1_q = s1.clone() if self.nq > 1 else s1
        if s2 is None:
            rows2, n2, conf2 = torch.zeros((1, sw), dtype=torch.int32, device=self.dev), 0, torch.zeros((1,), dtype=torch.float32, device=self.dev)
        else:
            rows2, n2, conf2 = s2.out_rows, s2.n_selected_dev, s2.selected_conf_dev
        rfs.append(devpipe.refine_device(s1_q, rows2, n2, conf2, d[f"isq{qi}"], d[f"sim{qi}"], plan.slice_points, cfg, self.names[qi],
                                         select_mode=1, comm=comm.handle))
    finals = [dist_mod.gather_final(comm, plan, rf, to_host and plan.rank == 0) for rf in rfs]
    ev()
    self.last_finals = finals
    return finals[0] if finals else None


GpuScene.step_sharded = _step_sharded


def _event():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def lift_algorithmic_bytes(g: GpuScene) -> int:
    """per frame 16*N (xyz) + depth bytes*H*W + M_f*H*W/8 + M_f*4*Wd, + 4*N*(1+Q) per scene.
    Depth is consumed as uint16 here (2 bytes per pixel), which is what the product path feeds."""
    n, hw = g.n, g.h * g.w
    wd = (n + 31) // 32
    depth_px = g.native[0] * g.native[1]            # the uint16 depth the kernel actually reads (native size)
    return g.f * (16 * n + 2 * depth_px) + g.n_masks * ((hw + 31) // 32 * 4 + 4 * wd) + 4 * n * (1 + g.nq)


class ClockSampler:
    QUERY_FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                    "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY_FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
            t0 = time.time()                       # nvidia-smi needs a moment before its first sample
            while time.time() - t0 < 5.0 and os.path.getsize(self.path) == 0:
                time.sleep(0.02)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                p = [x.strip() for x in line.split(",")]
                if len(p) < 6:
                    continue
                try:
                    sm.append(float(p[0]))
                    mx.append(float(p[1]))
                except ValueError:
                    continue
                for nm, v in zip(names, p[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(workload: str):
    """dram bytes per lift_kernel launch from the committed ncu --set full capture, if one exists."""
    p = os.path.join(ROOT, "profiles", "lift_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(workload)
        except Exception:
            return None
    return None


def run_gpu(args):
    from bff_b200 import _cabi, dist as dist_mod
    _cabi.lib()          # fail loudly when the CUDA library is missing
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (GPU arm) needs a CUDA device; there is no CPU fallback")
    ctx = dist_mod.init_from_env("nccl")
    rank, world = (ctx.rank, ctx.world_size) if ctx else (0, 1)
    local = ctx.local_rank if ctx else 0
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import torch.distributed as dist
    numa_cpus = dist_mod.bind_to_gpu_numa(local) if world > 1 else None     # staging buffers on the GPU's own socket
    # N > 1: scenes are the independent units of the path (the reference's outermost loop, :365), so the headline
    # run gives every rank its OWN c2-shaped scene and needs no data-path collective (weak scaling); the
    # frame-sharded variant of ONE scene (NCCL all-reduce + all-gather) is measured afterwards as `frame_sharded`.
    # (every rank fabricates the same seeded scene, so per-GPU work is identical by construction)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world > 1 and args.sharded_only:
        fs = frame_sharded_run(args, ctx, dev, flush, barrier)
        if rank == 0:
            emit({"metric": "point-view projections/sec (whole lifting+aggregation+refinement step)", "value": fs["value"], "unit": "point-views/s",
                  "n_gpus": world, "steps": fs["steps"], "warmup": args.warmup, "ms_per_step": fs["ms_per_step"], "higher_is_better": True,
                  "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                  "config": {"workload": f"{args.workload}: one scene sharded over {world} GPUs"}, "frame_sharded": fs})
        ctx.finalize()
        return
    inp = build_inputs(args.workload, 0, 1, dev)
    g = GpuScene(inp, dev, None)

    d = g.upload()
    torch.cuda.synchronize()
    final = None
    for _ in range(args.warmup):
        final = g.step(d)
        _ = g.step(g.upload())
    if args.pyprofile:
        import cProfile, pstats
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(args.steps):
            g.step(d)
        torch.cuda.synchronize()
        pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("cumulative").print_stats(45)
        pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(25)
        if ctx:
            ctx.finalize()
        return
    if args.profile:
        # for `ncu --profile-from-start off`: only these steps are captured; no number is reported
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for _ in range(args.steps):
            g.step(d)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        if ctx:
            ctx.finalize()
        return
    # ---- device-resident timing: K steps, each bracketed by events, L2 flushed in between (untimed). A step replays the
    # CUDA graph of the whole chain (one launch) and ends with the read-back of the final counts.
    graphed = g.graph_inputs(d)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    launches0 = _cabi.launch_count
    tot_ms = 0.0
    wall0 = time.time()
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0 = _event()
        final = g.step(d, to_host=False)
        e1 = _event()
        torch.cuda.synchronize()
        tot_ms += e0.elapsed_time(e1)
    launches = _cabi.launch_count - launches0
    barrier()
    wall_dev = time.time() - wall0
    clocks = sampler.stop() if rank == 0 else None
    # ---- the same steps kernel by kernel (no graph) with events at the stage boundaries and around the lifting launch:
    # the break-down and the roofline's launch time (a graph cannot be timed from the inside)
    lift_ms, stage_acc = 0.0, np.zeros(5)
    k_break = max(3, min(args.steps, 10))
    for _ in range(3):                                # the allocator pools of the ungraphed path (per query lane) warm up again
        g.step(d, [], to_host=False)
    torch.cuda.synchronize()
    for _ in range(k_break):
        flush.fill_(1)
        torch.cuda.synchronize()
        tm = []
        g.step(d, tm, to_host=False)
        torch.cuda.synchronize()
        stage_acc += np.asarray([tm[i].elapsed_time(tm[i + 1]) for i in range(len(tm) - 1)])
        lift_ms += g.lift_events[0].elapsed_time(g.lift_events[1])
    stage_acc *= args.steps / k_break
    lift_ms *= args.steps / k_break
    # ---- end to end: pinned host buffers -> device -> step -> final masks on the host
    e2e_ms = 0.0
    for _ in range(args.steps):
        flush.fill_(1)
        barrier()
        t0 = time.perf_counter()
        final = g.step(g.upload())
        torch.cuda.synchronize()
        e2e_ms += (time.perf_counter() - t0) * 1e3
    # ---- the same, as a scene loop is actually served: while scene i is processed, scene i+1's inputs are
    # already moving host -> device on a copy stream (every step still uploads all of its own inputs and
    # reads its result back inside the timed region; only the two are allowed to overlap)
    from bff_b200 import ops
    copy_stream = torch.cuda.Stream(device=dev, priority=-1 if args.upload.endswith('+hi') else 0)
    bufs = [{k: torch.empty_like(v, device=dev) for k, v in g.host.items()} for _ in range(2)]   # double buffer, no allocation in the loop
    # CUDA graphs and the copy engine do not mix here. Measured: a graph launch issued while the next scene's 184 MB move on
    # the H2D copy engine waits for that transfer -- the launch itself ships its command buffers over the same engine -- and
    # the loop went from 3.5 to 4.1 ms per scene at c2, from 7.9 to 42 ms at c4. So: with uploads on the copy engine (the
    # default for one query, where the loop is upload-bound anyway) the step is launched kernel by kernel; with uploads by
    # the SM copy kernel (--upload sm) the step replays its graph -- which measured slower still (c3: 13.9 vs 8.65 ms).
    done = [None, None]
    # The next scene's buffers move on the H2D copy engine (55 GB/s here, against 46-51 GB/s for an SM copy kernel that
    # shares SMs and the PCIe request queue with the running step). That only works because the running step never
    # touches that engine: its own small staging copies go through bff_upload (pipeline.cu upload_i32, ops.small_to_device);
    # a single cudaMemcpyAsync H2D inside the step would wait for the whole 184 MB transfer (measured before that change:
    # 4.8-5.2 ms per scene with the copy engine, 4.4 ms with the kernel; now 3.55 ms vs 3.95 ms).
    big = max(g.host, key=lambda k_: g.host[k_].numel() * g.host[k_].element_size())

    def _time_copy(fn):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); fn(); b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / 2
    t_ce = _time_copy(lambda: bufs[0][big].copy_(g.host[big], non_blocking=True))
    t_sm = _time_copy(lambda: ops.upload(bufs[0][big], g.host[big])) if g.host[big].numel() * g.host[big].element_size() >= 16 else float("inf")
    mode = args.upload.replace("+hi", "")   # "auto" | "sm" | "ce" | "ce:<MiB>" (copy engine in chunks, so that other streams' small copies interleave)
    if mode == "auto":
        # measured at c3 (20 queries): SM-kernel uploads + graph replay 13.9 ms per scene, copy-engine uploads + kernel-by-kernel
        # launches 8.65 ms -- the copy kernel shares SMs and the PCIe request queue with the step (r01 finding), so: the engine
        mode = "ce"
    use_sm_copy = mode == "sm"
    if use_sm_copy:
        for b_ in bufs:                               # one CUDA graph per buffer set (contents change, addresses do not)
            for k_, v_ in g.host.items():
                b_[k_].copy_(v_, non_blocking=True)
            g.graph_inputs(b_)
        torch.cuda.synchronize()
    chunk = int(float(mode.split(":")[1]) * (1 << 20)) if mode.startswith("ce:") else 0

    def upload_async(slot):
        with torch.cuda.stream(copy_stream):
            if done[slot] is not None:
                copy_stream.wait_event(done[slot])          # the step that last read this buffer has finished
            for k_, v_ in g.host.items():
                nbytes = v_.numel() * v_.element_size()
                if use_sm_copy and nbytes:
                    # --upload sm: EVERY buffer through the SM copy kernel. (Mixing is the worst choice: a copy-engine
                    # transfer queued in this stream behind a milliseconds-long copy kernel waits at the head of the
                    # engine's queue and holds back every other stream's copies.)
                    ops.upload(bufs[slot][k_], v_, copy_stream)
                elif chunk and nbytes > chunk:
                    src_b, dst_b = v_.view(-1).view(torch.uint8), bufs[slot][k_].view(-1).view(torch.uint8)
                    for a in range(0, nbytes, chunk):
                        dst_b[a:a + chunk].copy_(src_b[a:a + chunk], non_blocking=True)
                else:
                    bufs[slot][k_].copy_(v_, non_blocking=True)
            ev = torch.cuda.Event(enable_timing=diag)
            ev.record(copy_stream)
        return ev
    diag = bool(os.environ.get("BFF_E2E_DIAG"))         # dev: print the device timeline of uploads and steps (stderr)
    marks = []
    barrier()
    prof = None
    if os.environ.get("BFF_E2E_DIAG") == "2":
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    t0 = time.perf_counter()
    origin = _event() if diag else None
    nxt = upload_async(0)
    for i in range(args.steps):
        ev = nxt
        if i + 1 < args.steps:
            nxt = upload_async((i + 1) % 2)
        torch.cuda.current_stream().wait_event(ev)
        s_ev = _event() if diag else None
        tm_ = [] if diag else None
        final = g.step(bufs[i % 2], tm_)
        done[i % 2] = torch.cuda.Event(enable_timing=diag)
        done[i % 2].record()
        if diag:
            marks.append((ev, s_ev, done[i % 2], (time.perf_counter() - t0) * 1e3, tm_))
    torch.cuda.synchronize()
    e2e_pipe_ms = (time.perf_counter() - t0) * 1e3
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("tottime").print_stats(22)
    if diag:
        for i, (u, s0, s1, host_ms, tm_) in enumerate(marks):
            print(f"[e2e diag] scene {i}: upload done {origin.elapsed_time(u):7.3f}  step {origin.elapsed_time(s0):7.3f} -> {origin.elapsed_time(s1):7.3f}  host returned {host_ms:7.3f}"
                  f"  stages {[round(tm_[j].elapsed_time(tm_[j + 1]), 3) for j in range(len(tm_) - 1)]} lift {g.lift_events[0].elapsed_time(g.lift_events[1]):.3f}",
                  file=sys.stderr)
    d2h = sum(fn["ins"].numel() * fn["ins"].element_size() for fn in g.last_finals if isinstance(fn["ins"], torch.Tensor))
    # ---- IoU kernels alone: the aggregation-sized square call (tcgen05 int8) and a refinement-shaped call (200 x 27 rows, POPC)
    from bff_b200 import ops
    rows = g.lift(d, g.masks2d(d)).rows

    def _time_call(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0 = _event()
        for _ in range(reps):
            fn()
        e1 = _event()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    iou_ms = _time_call(lambda: ops.mask_iou(rows, want_iou=False))
    s1_rows = g.stage1_rows(d)
    few = rows[:27].contiguous()
    iou_small_ms = _time_call(lambda: ops.mask_iou(s1_rows, few, want_iou=False))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    int8_peak = 2.0 * float(peaks.get("bf16_tflops", 1672.3))             # TOP/s: int8 runs at twice the measured bf16 rate
    popc_peak = 148 * 16 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6    # word-ops/s: 16 POPC lanes per SM and clock
    n_r = rows.shape[0]
    iou_roofline = {
        "square": {"call": f"{n_r}x{n_r} rows x {g.n} points (tcgen05 kind::i8, popcounts fused)", "ms": iou_ms, "bound": "tensor",
                   "achieved": 2.0 * n_r * n_r * g.n / (iou_ms * 1e-3) / 1e12, "peak": int8_peak, "unit": "TOP/s (int8, full square counted)",
                   "frac": 2.0 * n_r * n_r * g.n / (iou_ms * 1e-3) / 1e12 / int8_peak,
                   "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops (burst)"},
        "few_columns": {"call": f"{s1_rows.shape[0]}x{few.shape[0]} rows x {g.n} points (POPC, one pass)", "ms": iou_small_ms, "bound": "popc",
                        "achieved": s1_rows.shape[0] * few.shape[0] * rows.shape[1] / (iou_small_ms * 1e-3) / 1e12, "peak": popc_peak / 1e12,
                        "unit": "T word-ops/s (AND+POPC)", "frac": s1_rows.shape[0] * few.shape[0] * rows.shape[1] / (iou_small_ms * 1e-3) / popc_peak,
                        "peak_source": "148 SMs x 16 POPC lanes/clk x sm_max_mhz"}}
    fs, fs_small = None, None
    if world > 1 and not args.no_frame_sharded:
        # The headline of an N-GPU run is ONE scene sharded over the GPUs (strong scaling). With the default workload that
        # scene is BASELINE configs[3] (c4s: 2M points, 2000 frames, 1296x968 masks over 640x480 depth, 500 stage-1 masks)
        # -- the configuration the reference metric names for 8 GPUs; the c2 scene of the 1-GPU line (0.6 ms on one GPU:
        # too small to shard, its ~45 small kernels and six all-reduces are latency) is measured too and kept as
        # `frame_sharded_c2`. An explicit --workload shards that workload instead.
        head = args.sharded_workload if args.workload == "c2" else args.workload
        if head != args.workload:
            fs_small = frame_sharded_run(args, ctx, dev, flush, barrier)
        fs = frame_sharded_run(args, ctx, dev, flush, barrier, workload=head)
    # ---- reduce over ranks (max time)
    t = torch.tensor([tot_ms, lift_ms, e2e_ms, e2e_pipe_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms, lift_ms, e2e_ms, e2e_pipe_ms = (float(x) for x in t.cpu())
    if rank != 0:
        if ctx:
            ctx.finalize()
        return
    k = args.steps
    pv = g.n * g.f * world
    ms_step, ms_lift, ms_e2e, ms_e2e_pipe = tot_ms / k, lift_ms / k, e2e_ms / k, e2e_pipe_ms / k
    peak, peak_src = measured_peak_hbm()
    alg = lift_algorithmic_bytes(g)
    achieved = alg / (ms_lift * 1e-3) / 1e9
    line = {
        "metric": "point-view projections/sec (whole lifting+aggregation+refinement step)",
        "value": pv / (ms_step * 1e-3), "unit": "point-views/s", "n_gpus": world, "steps": k, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: ScanNet200-shaped scene, {g.n} points, {g.f} frames/GPU {g.w}x{g.h}, "
                               f"{g.nq} quer{'y' if g.nq == 1 else 'ies'}, {g.n_masks} lifted 2D masks/GPU, {len(inp['stage1']['ins'])} stage-1 masks"
                               + (f", depth {g.native[1]}x{g.native[0]} resampled in the lift kernel" if g.native != (g.h, g.w) else ""),
                   "queries": g.nq,
                   "n_points": g.n, "frames_per_gpu": g.f, "lifted_masks_per_gpu": g.n_masks,
                   "parallelism": f"scene-parallel x{world}: one scene per GPU, no data-path collective" if world > 1 else "single GPU",
                   "l2": f"256 MB flush written between timed steps; per-step inputs are {g.h2d_bytes / 1e6:.0f} MB (> 126 MB L2)"},
        "stages_ms": dict(zip(["rle2d_to_bits", "lift_scene", "exchange", "aggregate_filter", "refine"],
                              (stage_acc / k).round(4).tolist())),
        "lift_point_views_per_s": pv / (ms_lift * 1e-3),
        "iou_pairs_per_s": (rows.shape[0] ** 2) / (iou_ms * 1e-3), "iou_call": f"{rows.shape[0]}x{rows.shape[0]} rows x {g.n} points, {iou_ms:.4f} ms",
        "iou_roofline": iou_roofline,
        "e2e": {"value": pv / (ms_e2e_pipe * 1e-3), "unit": "point-views/s", "h2d_bytes_per_step": int(g.h2d_bytes),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e_pipe,
                "how": "K consecutive scenes through the host API: pinned host inputs -> device -> step -> final masks on the "
                       "host; scene i+1's upload overlaps scene i's compute (copy stream); wall clock over the K steps",
                "serial_ms_per_step": ms_e2e, "serial_value": pv / (ms_e2e * 1e-3),
                "upload_engine": "SM copy kernel (bff_upload)" if use_sm_copy else f"copy engine (cudaMemcpyAsync{', %d MiB chunks' % (chunk >> 20) if chunk else ''})",
                "numa_bound_cpus_rank0": (len(numa_cpus) if numa_cpus else None),
                "upload_gbs_probe": {"copy_engine": round(g.host[big].numel() * g.host[big].element_size() / t_ce / 1e6, 1),
                                     "sm_kernel": round(g.host[big].numel() * g.host[big].element_size() / t_sm / 1e6, 1)}},
        "gpu_launches": int(launches), "step_as_cuda_graph": bool(graphed),
        "roofline": {"bound": "hbm", "kernel": "bff::lift_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": recorded_traffic(args.workload), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg, "ms_per_launch": ms_lift},
        "clocks": clocks,
    }
    if fs is not None:
        # N > 1: the measured headline is ONE scene sharded over the N GPUs (strong scaling: the scene and therefore the
        # total work is fixed); the N independent replicas measured above (one scene per GPU, no collective) are kept as
        # the secondary `replicas`
        line["replicas"] = {k_: line[k_] for k_ in ("value", "unit", "ms_per_step", "stages_ms", "e2e", "gpu_launches", "lift_point_views_per_s")}
        line["replicas"]["what"] = f"{world} independent {args.workload} scenes, one per GPU, no data-path collective (weak scaling)"
        pv1 = fs["n_points"] * fs["frames_total"]
        line.update({"value": fs["value"], "ms_per_step": fs["ms_per_step"], "scaling": "strong", "steps": fs["steps"],
                     "stages_ms": fs["stages_ms_rank0"], "lift_point_views_per_s": pv1 / (fs["lift_ms_per_launch_max"] * 1e-3)})
        line["config"]["workload"] = fs["workload_description"]
        line["config"]["n_points"], line["config"]["lifted_masks_per_gpu"] = fs["n_points"], fs["lifted_masks_total"] / world
        line["config"]["parallelism"] = (f"ONE {fs['workload']} scene on {world} GPUs: frame-sharded lifting ({fs['frames_total']} frames split), rows exchanged once "
                                         "(kernel-issued peer stores over NVLink, votes all-reduced), point-sharded aggregation + refinement")
        line["config"]["frames_per_gpu"] = fs["frames_total"] / world
        line["config"]["l2"] = "256 MB flush written between timed steps"
        line["e2e"] = {"value": pv1 / (fs["e2e_ms_per_step"] * 1e-3), "unit": "point-views/s", "h2d_bytes_per_step": fs["h2d_bytes_per_step"],
                       "d2h_bytes_per_step": fs["d2h_bytes_per_step"], "ms_per_step": fs["e2e_ms_per_step"],
                       "how": "every rank uploads its frames' inputs from pinned host memory, the sharded step, final masks on rank 0's host; wall clock, max over ranks"}
        alg_r = fs["lift_algorithmic_bytes_per_rank"]
        ach = alg_r / (fs["lift_ms_per_launch_max"] * 1e-3) / 1e9
        line["roofline"].update({"achieved": ach, "frac": ach / peak, "algorithmic_bytes_per_launch": alg_r, "ms_per_launch": fs["lift_ms_per_launch_max"],
                                 "traffic": None, "note": "per-rank launch over this rank's frames"})
        line["frame_sharded"] = fs
        if fs_small is not None:
            line["frame_sharded_c2"] = fs_small
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(inp, args.cpu_baseline_budget)
    emit(line)
    if ctx:
        ctx.finalize()


def frame_sharded_run(args, ctx, dev, flush, barrier, workload=None, steps=None):
    """STRONG scaling of ONE scene over the G ranks: rank r lifts its block of the scene's frames, one exchange (grouped
    NCCL send/recv of row slices + int32 vote all-reduce), then the point-sharded device-resident pipelines (int32
    all-reduces of the intersection matrices inside). Rank 0 also runs the same scene alone (the 1-GPU step of this
    very run) for the speed-up, and the final masks of both are compared bit for bit. Max over ranks."""
    import torch.distributed as dist
    from bff_b200 import dist as dist_mod
    workload = workload or args.workload
    rank, world = ctx.rank, ctx.world_size
    inp = build_inputs(workload, rank, world, dev, strong=True)
    g = GpuScene(inp, dev, ctx)
    # what every rank has to know about the others' rows (host metadata, once per scene): per query, the confidence, label,
    # launch position and FRAME of every lifted row. The scene's row order (the reference's file order: frames ascending,
    # masks in order) is then the stable sort of the rank blocks by frame.
    frame_of_row = [[int(r["frame_id"][:-4]) for r in recs for _ in r["segmented_frame_masks"]] for recs in inp["records_by_query"]]
    gathered = [None] * world
    dist.all_gather_object(gathered, (g.n_masks, [[float(c) for c in cq] for cq in g.conf_q], g.labels_q, g.row_index, g.f, frame_of_row))
    g.rows_per_rank = [x[0] for x in gathered]
    g.f_total = int(sum(x[4] for x in gathered))
    base = np.concatenate([[0], np.cumsum(g.rows_per_rank)])
    g.all_conf_dev, g.all_labels, g.all_index_dev32 = [], [], []
    for qi in range(g.nq):
        conf = [c for x in gathered for c in x[1][qi]]
        labels = [l for x in gathered for l in x[2][qi]]
        pos = [int(base[r]) + i for r, x in enumerate(gathered) for i in x[3][qi]]
        frame = [f for x in gathered for f in x[5][qi]]
        order = sorted(range(len(frame)), key=lambda k: frame[k])          # stable: rows of one frame keep their order
        g.all_conf_dev.append(torch.tensor([conf[k] for k in order], dtype=torch.float32).to(dev))
        g.all_labels.append([labels[k] for k in order])
        idx = [pos[k] for k in order]
        g.all_index_dev32.append(None if idx == list(range(int(base[-1]))) else torch.tensor(idx, dtype=torch.int32, device=dev))
    g.row0 = int(base[rank])
    torch.cuda.synchronize()
    g.comm = dist_mod.Comm(ctx, dev.index or 0)
    g.peers = None
    if os.environ.get("BFF_PEER_ROWS", "1") != "0" and world <= 8:
        try:
            g.peers = dist_mod.PeerRows(ctx, int(base[-1]), g.plan.slice_words, dev)
        except Exception as e:                      # IPC unavailable: the NCCL send/recv exchange stays
            if rank == 0:
                print(f"[bench] fused peer exchange unavailable ({e}); using the NCCL row exchange", file=sys.stderr)
    d = g.upload()
    for _ in range(3):
        final = g.step(d)
    steps = steps or max(1, args.steps)
    tot, stage = 0.0, np.zeros(5)
    barrier()
    for _ in range(steps):
        flush.fill_(1)
        barrier()
        tm = []
        final = g.step(d, tm, to_host=False)
        torch.cuda.synchronize()
        stage += np.asarray([tm[i].elapsed_time(tm[i + 1]) for i in range(len(tm) - 1)])
        tot += tm[0].elapsed_time(tm[-1])
    finals_sharded = list(g.last_finals)
    # ---- end to end: every rank uploads ITS frames' inputs from pinned host memory, the step, final masks to rank 0's host
    e2e = 0.0
    for _ in range(steps):
        flush.fill_(1)
        barrier()
        t0 = time.perf_counter()
        g.step(g.upload(), to_host=True)
        torch.cuda.synchronize()
        e2e += (time.perf_counter() - t0) * 1e3
    d2h = sum(fn["ins"].numel() * fn["ins"].element_size() for fn in g.last_finals if isinstance(fn["ins"], torch.Tensor)) if rank == 0 else 0
    barrier()
    t = torch.tensor([tot, e2e, float(g.h2d_bytes), g.lift_events[0].elapsed_time(g.lift_events[1])], dtype=torch.float64, device=dev)
    tmax = t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    ms, e2e_ms, h2d_total, lift_ms_max = float(tmax[0]) / steps, float(tmax[1]) / steps, float(t[2]), float(tmax[3])
    torch.cuda.synchronize()
    # ---- the same scene on ONE GPU (rank 0, the others wait): speed-up and bit-exact parity of the final masks
    single_ms, same = None, None
    if rank == 0:
        inp1 = build_inputs(workload, 0, 1, dev, strong=True)
        g1 = GpuScene(inp1, dev, None)
        d1 = g1.upload()
        for _ in range(3):
            g1.step(d1)
        g1.graph_inputs(d1)                           # the 1-GPU step as it is timed at N=1: one CUDA graph
        t1 = 0.0
        for _ in range(steps):
            flush.fill_(1)
            torch.cuda.synchronize()
            e0 = _event()
            g1.step(d1, to_host=False)
            e1 = _event()
            torch.cuda.synchronize()
            t1 += e0.elapsed_time(e1)
        single_ms = t1 / steps

        def same_final(a, b):
            if isinstance(a["ins"], list) or isinstance(b["ins"], list):
                return isinstance(a["ins"], list) and isinstance(b["ins"], list)
            return a["ins"].shape == b["ins"].shape and torch.equal(a["ins"].cpu(), b["ins"].cpu()) and \
                torch.equal(torch.as_tensor(a["conf"]), torch.as_tensor(b["conf"]))
        same = all(same_final(a, b) for a, b in zip(finals_sharded, g1.last_finals)) and len(finals_sharded) == len(g1.last_finals)
        if not same:
            raise AssertionError("frame-sharded final masks differ from the single-GPU result of the same scene")
    barrier()
    exchange_kind = "fused: lifting kernel stores row slices into the peers' buffers (CUDA IPC over NVLink)" if g.peers is not None else "grouped NCCL send/recv of row slices"
    if g.peers is not None:
        g.peers.close()
    g.comm.close()
    return {"value": g.n * g.f_total / (ms * 1e-3), "row_exchange": exchange_kind, "unit": "point-views/s", "ms_per_step": ms, "steps": steps, "scaling": "strong",
            "workload": workload, "n_points": g.n, "frames_total": g.f_total,
            "workload_description": f"{workload}: ONE ScanNet200-shaped scene, {g.n} points, {g.f_total} frames {g.w}x{g.h}, {g.nq} quer{'y' if g.nq == 1 else 'ies'}, "
                                    f"{int(sum(g.rows_per_rank))} lifted 2D masks, {len(inp['stage1']['ins'])} stage-1 masks"
                                    + (f", depth {g.native[1]}x{g.native[0]} resampled in the lift kernel" if g.native != (g.h, g.w) else ""), "lifted_masks_total": int(sum(g.rows_per_rank)),
            "stages_ms_rank0": dict(zip(["rle2d_to_bits", "lift_scene", "exchange", "aggregate_filter", "refine"],
                                        (stage / steps).round(4).tolist())),
            "e2e_ms_per_step": e2e_ms, "h2d_bytes_per_step": int(h2d_total), "d2h_bytes_per_step": int(d2h),
            "lift_ms_per_launch_max": lift_ms_max, "lift_algorithmic_bytes_per_rank": lift_algorithmic_bytes(g),
            "queries": g.nq, "single_gpu_ms_same_scene": single_ms, "speedup_vs_single_gpu": (single_ms / ms) if single_ms else None,
            "final_masks_identical_to_single_gpu": same,
            "collectives": "int32 all-reduce [votes], grouped send/recv [row slices], int32 all-reduces [intersections, popcounts] "
                           "inside the point-sharded pipelines; own NCCL communicator (bff_comm)"}


# ----------------------------------------------------------------------------- CPU baseline (oracle port)
# ----------------------------------------------------------------------------- config 5: the 312-scene sweep
SWEEP_SCENES, SWEEP_POOL = 312, 5


def sweep_plan(world: int, n_scenes: int = SWEEP_SCENES, n_pool: int = SWEEP_POOL, seed: int = 5):
    """BASELINE.json configs[4]: a val-set sweep of `n_scenes` ScanNet200-val-shaped scenes, scene-parallel over the ranks.
    Scene shapes are seeded draws over the ranges SURVEY F8 gives for the val set (40k-420k points, 58-543 frames, log-
    uniform). Fabricating 312 distinct scenes would take longer than the whole bench, so the shapes are bucketed by cost
    (points x frames) into `n_pool` quantile classes and every scene is served by its class's fabricated scene (the class's
    mean shape); each of the 312 passes still uploads all of its inputs from the host and brings its masks back.
    Returns (pool shapes [(points, frames)], class of every scene, per-rank scene lists by schedule.lpt_assign)."""
    from bff_b200 import schedule
    rng = np.random.default_rng(seed)
    pts = np.exp(rng.uniform(np.log(40_000), np.log(420_000), n_scenes))
    frs = np.exp(rng.uniform(np.log(58), np.log(543), n_scenes))
    cost = pts * frs
    order = np.argsort(cost, kind="stable")
    cls = np.empty(n_scenes, np.int64)
    pool = []
    for k, idx in enumerate(np.array_split(order, n_pool)):
        cls[idx] = k
        pool.append((int(round(pts[idx].mean() / 1000.0)) * 1000, int(round(frs[idx].mean()))))
    costs = [float(pool[c][0] * pool[c][1]) for c in cls]
    return pool, cls.tolist(), schedule.lpt_assign(costs, world)


def run_sweep(args):
    """`--workload c5`: every rank serves its LPT share of the 312 scenes end to end (pinned host buffers -> device ->
    whole step -> final masks on the host), scene i+1's upload overlapping scene i's step. No data-path collective: scenes
    are the reference's outermost loop (projection_2d_to_3d.py:365). value = point-views of ALL scenes / max-over-ranks time."""
    from bff_b200 import _cabi, dist as dist_mod
    _cabi.lib()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (GPU arm) needs a CUDA device; there is no CPU fallback")
    ctx = dist_mod.init_from_env("nccl")
    rank, world = (ctx.rank, ctx.world_size) if ctx else (0, 1)
    local = ctx.local_rank if ctx else 0
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    import torch.distributed as dist
    numa_cpus = dist_mod.bind_to_gpu_numa(local) if world > 1 else None
    pool, cls, shares = sweep_plan(world)
    mine = shares[rank]
    scenes = {}
    for c in sorted({cls[i] for i in mine}):
        n_pts, n_fr = pool[c]
        WORKLOADS[f"c5p{c}"] = dict(n_points=n_pts, n_frames=n_fr, width=640, height=480, n_stage1=max(20, min(200, n_pts // 1250)), n_boxes=14)
        inp = build_inputs(f"c5p{c}", 0, 1, dev, seed=5000 + c)
        g = GpuScene(inp, dev, None)
        bufs = [{k: torch.empty_like(v, device=dev) for k, v in g.host.items()} for _ in range(2)]
        scenes[c] = dict(g=g, bufs=bufs, done=[None, None], turn=0)
    copy_stream = torch.cuda.Stream(device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def upload_async(c):
        sc = scenes[c]
        slot = sc["turn"]
        sc["turn"] ^= 1
        with torch.cuda.stream(copy_stream):
            if sc["done"][slot] is not None:
                copy_stream.wait_event(sc["done"][slot])
            for k_, v_ in sc["g"].host.items():
                sc["bufs"][slot][k_].copy_(v_, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return slot, ev

    def serve(order):
        n_masks_out, d2h = 0, 0
        nxt = upload_async(cls[order[0]]) if order else None
        for j, i in enumerate(order):
            c = cls[i]
            slot, ev = nxt
            if j + 1 < len(order):
                nxt = upload_async(cls[order[j + 1]])
            torch.cuda.current_stream().wait_event(ev)
            sc = scenes[c]
            final = sc["g"].step(sc["bufs"][slot])
            sc["done"][slot] = torch.cuda.Event()
            sc["done"][slot].record()
            if isinstance(final["ins"], torch.Tensor):
                n_masks_out += final["ins"].shape[0]
                d2h += final["ins"].numel() * final["ins"].element_size()
        torch.cuda.synchronize()
        return n_masks_out, d2h

    for c in scenes:                                   # warm-up: every class of scene W times (workspaces, allocator pools)
        serve([next(i for i in mine if cls[i] == c)] * max(3, args.warmup))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = _cabi.launch_count
    barrier()
    t0 = time.perf_counter()
    n_out, d2h = serve(mine)
    ms = (time.perf_counter() - t0) * 1e3
    launches = _cabi.launch_count - launches0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    pv_mine = float(sum(pool[cls[i]][0] * pool[cls[i]][1] for i in mine))
    h2d_mine = float(sum(scenes[cls[i]]["g"].h2d_bytes for i in mine))
    t = torch.tensor([ms, -ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([pv_mine, h2d_mine, float(d2h), float(n_out), float(launches), float(len(mine))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_max, ms_min = float(t[0]), -float(t[1])
    pv, h2d, d2h_all, n_out_all, launches_all, n_served = (float(x) for x in tot.cpu())
    if rank == 0:
        assert int(n_served) == SWEEP_SCENES
        emit({"metric": "point-view projections/sec (whole lifting+aggregation+refinement step)", "value": pv / (ms_max * 1e-3),
              "unit": "point-views/s", "n_gpus": world, "steps": SWEEP_SCENES, "warmup": max(3, args.warmup), "ms_per_step": ms_max / SWEEP_SCENES * world,
              "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
              "config": {"workload": f"c5: sweep of {SWEEP_SCENES} ScanNet200-val-shaped scenes (40k-420k points, 58-543 frames 640x480, 1 query), "
                                     f"{SWEEP_POOL} distinct fabricated scenes serve the {SWEEP_SCENES} passes by size class: "
                                     + ", ".join(f"{p_}x{f_}" for p_, f_ in pool),
                         "parallelism": f"scene-parallel: longest-processing-time-first over {world} GPU(s) (bff_b200/schedule.py), no data-path collective",
                         "scenes_per_rank": [len(s_) for s_ in shares], "l2": "every pass re-uploads its scene (36-390 MB > L2 for all but the smallest class)"},
              "sweep_ms": ms_max, "sweep_ms_fastest_rank": ms_min, "scenes_per_s": SWEEP_SCENES / (ms_max * 1e-3), "final_masks": int(n_out_all),
              "e2e": {"value": pv / (ms_max * 1e-3), "unit": "point-views/s", "h2d_bytes_per_step": int(h2d / SWEEP_SCENES), "d2h_bytes_per_step": int(d2h_all / SWEEP_SCENES),
                      "how": "the sweep IS end to end: every pass copies its inputs from pinned host memory and its final masks back; wall clock, max over ranks",
                      "h2d_gbs_per_gpu": h2d / world / (ms_max * 1e-3) / 1e9, "numa_bound_cpus_rank0": (len(numa_cpus) if numa_cpus else None)},
              "gpu_launches": int(launches_all), "clocks": clocks})
    if ctx:
        ctx.finalize()


def cpu_frames_of(inp, n_frames):
    """First n_frames of the rank's block as the reference's per-frame inputs (dense float32 masks / depth at mask
    resolution, cv2.resize like :436 when the depth PNGs are smaller). Returns (per-query mask frames, view frames)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import reference_port as RP
    from bff_b200 import scene_io
    hgt, wid = inp["spec"].height, inp["spec"].width
    view_frames = []
    for i in range(min(n_frames, len(inp["frame_ids"]))):
        view_frames.append({"pose": inp["poses"][i], "depth": scene_io.depth_to_metres(inp["depth"][i], wid, hgt)})
    per_query = []
    for recs in inp["records_by_query"]:
        by_frame = {r["frame_id"][:-4]: r for r in recs}
        mask_frames = []
        for i, fid in enumerate(inp["frame_ids"][:n_frames]):
            if fid in by_frame:
                r = by_frame[fid]
                masks = np.stack([RP.rle_to_mask(m).reshape(hgt, wid) for m in r["segmented_frame_masks"]]).astype(np.float32)
                mask_frames.append(dict(view_frames[i], masks=masks, conf=r["confidences"], labels=r["labels"]))
        per_query.append(mask_frames)
    return per_query, view_frames


def cpu_step(inp, mask_frames_by_query, view_frames):
    """The reference's path for one scene on the CPU (numpy/OpenBLAS + torch CPU, oracle/reference_port.py): one full
    run per text query, each projecting every frame again, exactly like one reference process per --cls."""
    import reference_port as RP
    from bff_b200.refine import STAGE1_CLASS_NAMES
    cfg = inp["cfg"]
    out = []
    for name, mask_frames in zip(inp["queries"], mask_frames_by_query):
        res2 = RP.stage2_scene(inp["xyz"], inp["K"], mask_frames, view_frames, cfg)
        s1 = torch.from_numpy(np.stack([RP.rle_to_mask(r) for r in inp["stage1"]["ins"]]))
        labels = [STAGE1_CLASS_NAMES[int(i)] for i in inp["stage1"]["final_class"]]
        p1 = RP.refine_pass1(s1, labels, res2, name, cfg["stage1_iou_thres"], fake_similarity)
        thr = RP.sim_threshold([p1["sims"]], cfg["refinment_sim_percentile"]) if len(p1["sims"]) else 0.0
        out.append(RP.refine_pass2(p1, name, cfg["refiment_iou_thres"], thr))
    return out


def _set_cpu_threads(k: int):
    """Threads of torch's intra-op pool and of the BLAS / OpenMP pools numpy and OpenCV use (threadpoolctl)."""
    torch.set_num_threads(max(1, int(k)))
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=max(1, int(k)))
    except Exception:
        pass


def cpu_arm_plan(inp, budget_s: float, runs: int):
    """How the reference's CPU path is timed on this host within `budget_s` seconds for `runs` passes: a probe over the
    first frames at 1 thread and at all cores (the path is thread-setting sensitive: numpy's K<=4 dgemms and the dense
    float32 sgemms pull in opposite directions), the faster setting is kept, and the number of frames per pass follows
    from the probe's per-frame cost. Returns (threads, n_frames, note)."""
    cores = os.cpu_count() or 1
    f_all = inp["depth"].shape[0]
    n_probe = min(f_all, max(6, 24 // max(len(inp["queries"]), 1)))
    mask_frames, view_frames = cpu_frames_of(inp, n_probe)
    probe = {}
    for k in sorted({1, cores}):
        _set_cpu_threads(k)
        t0 = time.perf_counter()
        cpu_step(inp, mask_frames, view_frames)
        probe[k] = time.perf_counter() - t0
    threads = min(probe, key=probe.get)
    per_frame = probe[threads] / n_probe
    n_frames = int(max(n_probe, min(f_all, (budget_s / max(runs, 1)) / max(per_frame, 1e-6))))
    note = "thread sweep on the first %d frames: %s; kept %d thread(s)" % (n_probe, ", ".join(f"{k} thr {v:.2f} s" for k, v in probe.items()), threads)
    return threads, n_frames, note


def _lifted_masks_in(inp, n_frames):
    ids = set(inp["frame_ids"][:n_frames])
    return [sum(len(r["segmented_frame_masks"]) for r in recs if r["frame_id"][:-4] in ids) for recs in inp["records_by_query"]]


def cpu_baseline(inp, budget_s: float = 20.0):
    threads, n_frames, note = cpu_arm_plan(inp, budget_s, 1)
    _set_cpu_threads(threads)
    mask_frames, view_frames = cpu_frames_of(inp, n_frames)
    t0 = time.perf_counter()
    cpu_step(inp, mask_frames, view_frames)
    dt = time.perf_counter() - t0
    n = inp["spec"].n_points
    f_all = inp["depth"].shape[0]
    return {"value": n * n_frames / dt, "unit": "point-views/s", "cores": threads, "host_cores": os.cpu_count() or 1, "kind": "port",
            "frames": n_frames, "frames_full": f_all, "lifted_masks": _lifted_masks_in(inp, n_frames), "same_config": n_frames == f_all,
            "sample": f"first {n_frames} of {f_all} frames of the same scene ({n} points), whole path (lift + viewed pass + aggregate + "
                      f"filter + refine), {dt:.2f} s, oracle/reference_port.py (numpy/OpenBLAS + torch CPU like the reference); {note}"}


# ----------------------------------------------------------------------------- files -> files (BASELINE configs[0])
def _install_fake_clip():
    """CLIP text similarity is an INPUT of refinement (refinement.py:93-117) and the model is not in this image: a
    deterministic stand-in module named `clip` so that tools/refinement.py's own code path (ClipTextSimilarity) runs."""
    import hashlib
    import types

    def vec(text):
        g = torch.Generator().manual_seed(int.from_bytes(hashlib.sha1(text.encode()).digest()[:4], "little"))
        return torch.randn(1, 32, generator=g) + 1.0

    class Tok:
        def __init__(self, texts):
            self.texts = list(texts)

        def to(self, *_a, **_k):
            return self

    class Model:
        def encode_text(self, tok):
            return torch.cat([vec(t) for t in tok.texts], dim=0)
    m = types.ModuleType("clip")
    m.load = lambda name, device=None: (Model(), None)
    m.tokenize = lambda texts: Tok(texts)
    sys.modules["clip"] = m

    def sim(query, label):
        a, b = vec(query), vec(label)
        s_ = a @ b.T
        return (s_ / (a.norm(dim=-1, keepdim=True) * b.norm(dim=-1, keepdim=True).T)).item()
    return sim


def run_cli_e2e(args):
    """--cli-e2e: BASELINE configs[0] as a user runs it -- an on-disk tree in the reference's formats (npy cloud, pose /
    intrinsic text, uint16 depth PNGs, RLE .pth masks), `tools/projection_2d_to_3d.py` then `tools/refinement.py` of this
    repo in-process, files to files, wall clock; next to it the reference's path (oracle port: per-frame imread + numpy
    projection + torch CPU aggregation/refinement) over the same tree on the host cores. Prints one JSON line."""
    import shutil
    import yaml
    from bff_b200 import refine, scene_io, stage2, synth
    dev = torch.device("cuda", 0)
    w = WORKLOADS[args.workload]
    root = tempfile.mkdtemp(prefix="bff_cli_")
    n_scenes = 3
    scenes = {}
    for i in range(n_scenes):
        spec = synth.SceneSpec(n_points=w["n_points"], n_frames=w["n_frames"], width=w["width"], height=w["height"],
                               depth_width=w.get("depth_width", 0), depth_height=w.get("depth_height", 0), n_boxes=w["n_boxes"],
                               n_stage1=w["n_stage1"], seed=3000 + i, n_queries=1, room=w.get("room", (6.0, 5.0, 2.8)))
        scenes[f"scene{i + 1:04d}_00"] = synth.make_scene(spec)
    cfg_path = synth.write_tree(root, scenes, [QUERY], device=dev)
    cfg = yaml.safe_load(open(cfg_path))
    sim = _install_fake_clip()
    cwd = os.getcwd()
    os.chdir(root)
    try:
        stage2.main(["--config", cfg_path, "--cls", QUERY])                    # warm-up pass (CUDA context, library load)
        shutil.rmtree(os.path.join(root, "checkpoints")); os.makedirs(os.path.join(root, "checkpoints"))
        t0 = time.perf_counter()
        stage2.main(["--config", cfg_path, "--cls", QUERY])
        t1 = time.perf_counter()
        refine.main(["--config", cfg_path, "--cls", QUERY])
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        ours = {sid: scene_io.torch_load(os.path.join(cfg["final_output_dir"], QUERY, f"{sid}.pth")) for sid in scenes}
        # ---- the reference's path over the same files (oracle port; the one CPU leg)
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from helpers import port_run_class
        import helpers
        helpers.fake_clip_sim = sim
        torch.set_num_threads(os.cpu_count() or 1)
        t3 = time.perf_counter()
        _, finals = port_run_class(cfg, QUERY)
        t4 = time.perf_counter()
    finally:
        os.chdir(cwd)
        shutil.rmtree(root, ignore_errors=True)
    same = all(isinstance(ours[s]["ins"], torch.Tensor) and isinstance(finals[s]["ins"], torch.Tensor)
               and torch.equal(ours[s]["ins"].cpu(), finals[s]["ins"]) for s in scenes)
    pv = n_scenes * w["n_points"] * w["n_frames"]
    emit({"metric": "point-view projections/sec, files to files (both CLIs, disk in, disk out)", "value": pv / (t2 - t0),
          "unit": "point-views/s", "n_gpus": 1, "higher_is_better": True, "data": "synthetic", "dtype": "f64",
          "config": {"workload": f"{args.workload}: {n_scenes} scenes x {w['n_points']} points x {w['n_frames']} frames "
                                 f"{w['width']}x{w['height']}, 1 query, on-disk tree in the reference's formats"},
          "seconds": {"lifting_cli": t1 - t0, "refinement_cli": t2 - t1, "per_scene": (t2 - t0) / n_scenes},
          "cpu_baseline": {"value": pv / (t4 - t3), "unit": "point-views/s", "cores": os.cpu_count(), "kind": "port",
                           "sample": f"the same {n_scenes} scenes, whole path from the same files, {t4 - t3:.1f} s"},
          "masks_identical_to_cpu_path": bool(same)})


def run_reference(args):
    """--impl reference: the reference's CPU path on the host cores, same metric and config. The unmodified reference
    scripts cannot travel to the GPU box (/root/reference exists only in the build container, where
    oracle/time_reference_cli.py times them: profiles/r02_reference_cli_c1_container.json), so this arm times
    oracle/reference_port.py -- the same numpy / OpenBLAS / torch-CPU calls in the same order, pinned to the reference's
    outputs by tests/test_port_e2e.py -- at the faster of {1 thread, all cores}, on as many frames of the scene as fit
    the time budget (all of them for small K: then `same_config` is true)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")   # input fabrication only
    inp = build_inputs(args.workload, 0, 1, dev)
    warm = min(args.warmup, 1)
    threads, n_frames, note = cpu_arm_plan(inp, args.cpu_budget, args.steps + warm)
    _set_cpu_threads(threads)
    mask_frames, view_frames = cpu_frames_of(inp, n_frames)
    for _ in range(warm):
        cpu_step(inp, [m[:2] for m in mask_frames], view_frames[:2])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(inp, mask_frames, view_frames)
    dt = (time.perf_counter() - t0) / args.steps
    n = inp["spec"].n_points
    f_all = inp["depth"].shape[0]
    v = n * n_frames / dt
    sample = (f"each step = first {n_frames} of {f_all} frames of the scene ({n} points), whole path; oracle/reference_port.py "
              f"(numpy/OpenBLAS + torch CPU); {note}")
    emit({
        "impl": "reference", "metric": "point-view projections/sec (whole lifting+aggregation+refinement step)",
        "value": v, "unit": "point-views/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: ScanNet200-shaped scene, {n} points, {f_all} frames, {len(inp['queries'])} quer{'y' if len(inp['queries']) == 1 else 'ies'}",
                   "frames": n_frames, "frames_full": f_all, "lifted_masks": _lifted_masks_in(inp, n_frames), "same_config": n_frames == f_all,
                   "n_points": n},
        "cpu_baseline": {"value": v, "unit": "point-views/s", "cores": threads, "host_cores": os.cpu_count() or 1, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "point-views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


_REAL_STDOUT = None


def emit(line: dict):
    """The ONE JSON line goes to the process's original stdout; everything else that libraries print to
    fd 1 (NCCL's version banner, for instance) was diverted to stderr in main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + ["c5"])
    ap.add_argument("--cpu-budget", type=float, default=150.0, help="--impl reference: seconds of CPU work for all steps together (sizes the frame sample)")
    ap.add_argument("--cpu-baseline-budget", type=float, default=20.0, help="our arm at N=1: seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-frame-sharded", action="store_true", help="N>1: skip the frame-sharded (strong scaling) measurement")
    ap.add_argument("--sharded-workload", default="c4s", choices=sorted(WORKLOADS), help="N>1 with the default workload: the scene that is sharded for the headline line")
    ap.add_argument("--sharded-only", action="store_true", help="N>1: only the sharded scene (with its single-GPU cross-check), no replicas")
    ap.add_argument("--pyprofile", action="store_true", help="cProfile K steps of the host path (stderr) and exit")
    ap.add_argument("--cli-e2e", action="store_true", help="files-to-files run of both drop-in CLIs on an on-disk tree (configs[0]) and exit")
    ap.add_argument("--upload", default="auto", help="pipelined e2e: how scene i+1's buffers move while scene i runs: ce (copy engine) | sm (bff_upload kernel) | ce:<MiB>")
    ap.add_argument("--profile", action="store_true", help="run K steps between cudaProfilerStart/Stop and exit (for ncu)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.cli_e2e:
        run_cli_e2e(args)
    elif args.impl == "reference":
        run_reference(args)
    elif args.workload == "c5":
        run_sweep(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
