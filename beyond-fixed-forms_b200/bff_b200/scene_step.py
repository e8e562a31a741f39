"""One scene, all its queries, one enqueue: lift -> aggregate -> point filter -> overlap resolution -> refinement.

This is the whole hot path of the reference for one (scene, queries) unit -- the frame loops and everything after them in
tools/projection_2d_to_3d.py:413-634, then both passes of tools/refinement.py:182-426 -- expressed as a sequence of device
work without a single host round trip (devpipe.py takes every decision on the GPU). Because nothing in it waits for the
host, the sequence can be recorded ONCE as a CUDA graph and replayed per scene: one launch instead of ~60, no Python or
driver latency between kernels, and the queries' chains run concurrently on their own streams inside the graph.

    step = SceneStep(...static description of the scene shape...)
    finals = step.run(d)           # d: dict of device tensors holding the scene's inputs
    step.capture(d); step.run(d)   # the same, replaying a CUDA graph bound to d's buffers
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import _cabi, devpipe, ops


class SceneStep:
    def __init__(self, *, n_points: int, height: int, width: int, intrinsics, native_size, names: Sequence[str],
                 labels_q: Sequence[Sequence[str]], cfg: dict, n_stage1: int, n_masks: int, n_frames: int,
                 row_index: Optional[Sequence[torch.Tensor]], device, lanes: int = 4, depth_range=(0.0, 65.535)):
        """names: the queries; labels_q[q]: labels of query q's lifted rows in file order; row_index[q]: int32 CUDA launch
        positions of query q's rows (None for a single query: all rows, in order)."""
        self.n, self.h, self.w, self.K, self.native = int(n_points), int(height), int(width), intrinsics, native_size
        self.names, self.labels_q, self.cfg = list(names), [list(l) for l in labels_q], cfg
        self.n1, self.n_masks, self.f = int(n_stage1), int(n_masks), int(n_frames)
        self.row_index, self.dev, self.depth_range = row_index, device, depth_range
        self.nq = len(self.names)
        self.lanes = [torch.cuda.Stream(device=device) for _ in range(min(max(int(lanes), 1), self.nq))] if self.nq > 1 else []
        self.side = torch.cuda.Stream(device=device)      # the stage-1 decode: needed only by the refinement, runs beside the lifting
        self._graphs: dict = {}
        self.timing_events = None

    # ------------------------------------------------------------------ enqueue-only part
    def _query_chain(self, qi: int, d: dict, out, s1: torch.Tensor, lane: int, mark_mid=None, join=None):
        rows_q = out.rows if self.row_index is None else ops.rows_gather(out.rows, self.row_index[qi])
        wd = s1.shape[1]
        if rows_q.shape[0] == 0:                       # nothing lifted for this query (:465-478): known on the host
            s2 = None
            rows2, n2 = torch.zeros((1, wd), dtype=torch.int32, device=self.dev), 0
            conf2 = torch.zeros((1,), dtype=torch.float32, device=self.dev)
        else:
            s2 = devpipe.stage2_device(rows_q, d[f"conf{qi}"], self.labels_q[qi], out.masked_counts[qi], out.viewed_counts, self.n, self.cfg,
                                       mc_cap=rows_q.shape[0], vc_cap=self.f, lane=lane)
            rows2, n2, conf2 = s2.out_rows, s2.n_selected_dev, s2.selected_conf_dev
        if mark_mid is not None:
            mark_mid()
        if join is not None:
            join()
        s1_q = s1.clone() if self.nq > 1 else s1     # pass 1 ORs matched stage-1 rows in place (refinement.py:248)
        rf = devpipe.refine_device(s1_q, rows2, n2, conf2, d[f"isq{qi}"], d[f"sim{qi}"], self.n, self.cfg, self.names[qi], select_mode=1, lane=lane)
        rf.enqueue_unpack()
        return s2, rf

    def enqueue(self, d: dict, marks=None):
        """Everything of the scene, enqueued on the current stream (and the query lanes forked from / joined into it).
        `marks`: optional callable invoked at the stage boundaries (the caller records timing events there)."""
        mark = marks or (lambda: None)
        mark()
        # the stage-1 masks depend on nothing and are needed only by the refinement: decoded on a side stream, beside the lifting
        main = torch.cuda.current_stream(self.dev)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            s1 = ops.rle_pairs_to_bits_dev(d["s1_counts"], d["s1_pair_off"], self.n)
        joined = [False]

        def join():
            if not joined[0]:
                torch.cuda.current_stream(self.dev).wait_stream(self.side)
                joined[0] = True
        masks2d = ops.rle_pairs_to_bits_dev(d["counts"], d["pair_off"], self.h * self.w) if self.n_masks else None
        mark()
        out = ops.lift_scene(d["xyz"], d["pinv"], self.K, d["depth"], self.h, self.w, masks2d=masks2d,
                             mask_offset=d["offsets"] if self.n_masks else None, mask_query=d.get("mask_query") if self.n_masks else None,
                             n_queries=self.nq, frame_viewed=d["viewed"], depth_thresh=0.08, depth_scale=1000.0, depth_range=self.depth_range,
                             native_size=self.native, timing_events=self.timing_events)
        mark()
        mark()                                         # (a single GPU has no exchange stage: an empty interval)
        chains: List = [None] * self.nq
        if self.lanes:
            join()                                     # before the fork: every lane then sees the decoded stage-1 rows
            for k, lane in enumerate(self.lanes):
                lane.wait_stream(main)
            for qi in range(self.nq):
                k = qi % len(self.lanes)
                with torch.cuda.stream(self.lanes[k]):
                    chains[qi] = self._query_chain(qi, d, out, s1, k + 1)
            for lane in self.lanes:
                main.wait_stream(lane)
            mark()
        else:
            for qi in range(self.nq):
                chains[qi] = self._query_chain(qi, d, out, s1, 0, mark_mid=mark if self.nq == 1 else None, join=join)
            if self.nq != 1:
                mark()
        return {"s1": s1, "out": out, "chains": chains, "masks2d": masks2d}

    # ------------------------------------------------------------------ the one synchronisation
    def finish(self, state: dict, to_host: bool = True):
        finals = [rf.final(to_host) for (_, rf) in state["chains"]]
        return finals

    # ------------------------------------------------------------------ CUDA graph
    @staticmethod
    def _key(d: dict):
        return tuple(sorted((k, v.data_ptr()) for k, v in d.items()))

    def capture(self, d: dict):
        """Record enqueue(d) as a CUDA graph bound to the buffers of `d` (their CONTENTS may change between replays).
        A plain run comes first: workspaces, side streams and kernel attributes must exist before the capture."""
        key = self._key(d)
        if key in self._graphs:
            return
        ev = self.timing_events
        self.timing_events = None                     # events recorded inside a capture cannot be timed
        try:
            self.finish(self.enqueue(d), to_host=False)
            torch.cuda.synchronize(self.dev)
            launches0 = _cabi.launch_count
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                state = self.enqueue(d)
            self._graphs[key] = (g, state, _cabi.launch_count - launches0)
        finally:
            self.timing_events = ev

    def has_graph(self, d: dict) -> bool:
        return bool(self._graphs) and self._key(d) in self._graphs

    def run(self, d: dict, to_host: bool = True, marks=None):
        """The whole scene. Replays the graph captured for `d` when there is one, else enqueues kernel by kernel.
        marks (stage boundaries, ungraphed runs only): start | RLE decode | lifting | (exchange) | aggregate + filter |
        refinement + read-back."""
        hit = self._graphs.get(self._key(d)) if self._graphs else None
        if hit is not None and marks is None:
            g, state, launches = hit
            g.replay()
            _cabi.launch_count += launches
            for (_, rf) in state["chains"]:
                rf._host = None                       # same buffers, new contents
            return self.finish(state, to_host), state
        state = self.enqueue(d, marks)
        finals = self.finish(state, to_host)
        if marks is not None:
            marks()
        return finals, state
