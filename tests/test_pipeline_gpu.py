"""GPU: the fused C++ pipelines (bff_stage2_aggregate, bff_refine_match) give exactly the results of the
step-by-step Python/torch path (which the golden end-to-end tests pin to the reference), on scenes with
hundreds of lifted masks, both point-filter modes, and against the CPU oracle port."""
import os
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _run(inp, g, d, fused, cfg):
    from bff_b200 import ops, refine, stage2
    import bench
    stage2.FUSED = refine.FUSED = fused
    try:
        masks2d = g.masks2d(d)
        out = ops.lift_scene(d["xyz"], d["pinv"], inp["K"], d["depth"], g.h, g.w, masks2d=masks2d, mask_offset=d["offsets"],
                             frame_viewed=d["viewed"], depth_range=(0.0, 65.535))
        res2, _ = stage2.aggregate_and_filter(out.rows, g.conf_list, g.labels, out.masked_counts[0], out.viewed_counts, g.n, cfg,
                                              keep_rows=fused)
        s1 = g.stage1_rows(d)
        p1 = refine.pass1_scene(inp["stage1"], res2, bench.QUERY, cfg["stage1_iou_thres"], bench.fake_similarity, g.dev, s1_rows=s1)
        thr = refine.similarity_threshold([p1["sims"]], cfg["refinment_sim_percentile"]) if len(p1["sims"]) else 0.0
        return res2, p1, refine.pass2_scene(p1, bench.QUERY, cfg["refiment_iou_thres"], thr)
    finally:
        stage2.FUSED = refine.FUSED = True


def _same(a, b, what):
    assert list(a["final_class"]) == list(b["final_class"]), what
    if isinstance(a["ins"], list):
        assert isinstance(b["ins"], list) and len(a["ins"]) == len(b["ins"]) == 0, what
        return
    assert torch.equal(a["ins"], b["ins"]), what
    assert torch.equal(torch.as_tensor(a["conf"]), torch.as_tensor(b["conf"])), what


@pytest.mark.parametrize("occurrence,seed", [(False, 2000), (True, 2000), (False, 2011), (True, 2012), (False, 2013), (False, 2014)])
def test_fused_pipeline_equals_stepwise_and_port(cuda_dev, occurrence, seed):
    sys.path.insert(0, ROOT)
    import bench
    inp = bench.build_inputs("tiny", 0, 1, cuda_dev, seed=seed)
    cfg = dict(inp["cfg"])
    if occurrence:
        cfg.update(if_occurance_threshold=True, if_detected_ratio_threshold=False)
    inp["cfg"] = cfg
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    assert g.n_masks > 8
    a2, ap1, afin = _run(inp, g, d, True, cfg)
    b2, bp1, bfin = _run(inp, g, d, False, cfg)
    assert a2["ins"].shape[0] >= 1 and afin["ins"].shape[0] >= 1
    _same(a2, b2, "stage 2")
    _same(afin, bfin, "final")
    assert torch.equal(torch.as_tensor(ap1["ious"]), torch.as_tensor(bp1["ious"])) and ap1["sims"] == bp1["sims"]
    # and the dense CPU port of the reference (oracle) on the same inputs
    mask_frames, view_frames = bench.cpu_frames_of(inp, inp["depth"].shape[0])
    pfin = bench.cpu_step(inp, mask_frames, view_frames)[0]
    assert torch.equal(afin["ins"], pfin["ins"]) and list(afin["final_class"]) == list(pfin["final_class"])
    np.testing.assert_allclose(afin["conf"].numpy(), pfin["conf"].numpy(), rtol=1e-5)


def test_concurrent_queries_equal_sequential_loop_and_port(cuda_dev, monkeypatch):
    """Three text queries of one scene, depth PNGs at half the mask size (resampled in the lift kernel): the per-query
    chains on concurrent stream lanes with the decisions on the device (scene_step.py) must give what the host-decision
    batch gives, and both what the dense CPU port of the reference computes with one full run per query (cv2.resize on
    every frame)."""
    sys.path.insert(0, ROOT)
    import bench
    inp = bench.build_inputs("tiny3", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    assert g.nq == 3 and g.native == (120, 160) and (g.h, g.w) == (240, 320)
    d = g.upload()
    g.device_decisions = True
    g.step(d)
    pooled = list(g.last_finals)
    g.graph_inputs(d)                         # and replayed as one CUDA graph
    g.step(d)
    graphed = list(g.last_finals)
    g.device_decisions = False
    g.step(d)
    serial = list(g.last_finals)
    mask_frames, view_frames = bench.cpu_frames_of(inp, inp["depth"].shape[0])
    port = bench.cpu_step(inp, mask_frames, view_frames)
    assert sum(isinstance(f["ins"], torch.Tensor) and f["ins"].shape[0] >= 2 for f in serial) >= 2
    for qi in range(g.nq):
        _same(pooled[qi], serial[qi], f"lanes vs host batch, query {qi}")
        _same(graphed[qi], serial[qi], f"CUDA graph vs host batch, query {qi}")
        assert list(serial[qi]["final_class"]) == list(port[qi]["final_class"])
        if isinstance(serial[qi]["ins"], torch.Tensor):
            assert torch.equal(serial[qi]["ins"].cpu(), port[qi]["ins"])
            np.testing.assert_allclose(torch.as_tensor(serial[qi]["conf"]).cpu().numpy(), port[qi]["conf"].numpy(), rtol=1e-5)


def test_precomputed_pairs_give_the_same_aggregation(cuda_dev):
    """bff_stage2_aggregate with inter_pre / pop_pre (what dist.sharded_pairs assembles from the ranks' row blocks)."""
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops, stage2
    inp = bench.build_inputs("tiny", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    out = g.lift(d, g.masks2d(d))
    a, _ = stage2.aggregate_and_filter(out.rows, g.conf_list, g.labels, out.masked_counts[0], out.viewed_counts, g.n, inp["cfg"])
    top, bottom = out.rows[:11], out.rows[11:]                                # two "rank" blocks of the all-pairs matrix
    i0, p0, _, _ = ops.mask_iou(top, out.rows, want_iou=False)
    i1, p1, _, _ = ops.mask_iou(bottom, out.rows, want_iou=False)
    pairs = (torch.cat([i0, i1]), torch.cat([p0, p1]))
    b, _ = stage2.aggregate_and_filter(out.rows, g.conf_list, g.labels, out.masked_counts[0], out.viewed_counts, g.n, inp["cfg"], pairs=pairs)
    assert a["ins"].shape[0] >= 2
    _same(a, b, "precomputed pairs")


def test_batched_aggregation_equals_one_call_per_query(cuda_dev):
    """bff_stage2_aggregate_batch (all queries advance phase by phase, shared synchronisations) against one
    bff_stage2_aggregate call per query; with an empty query and one whose masks never reach min_aggragated_masks."""
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops, stage2
    inp = bench.build_inputs("tiny3", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    out = g.lift(d, g.masks2d(d))
    cfg = inp["cfg"]
    jobs = []
    for qi in range(g.nq):
        idx = g.row_index_dev[qi] if g.row_index_dev is not None else None
        jobs.append((out.rows if idx is None else out.rows[idx], g.conf_q[qi], g.labels_q[qi], out.masked_counts[qi], out.viewed_counts, g.n))
    jobs.append((None, [], [], None, None, g.n))                                  # nothing lifted for this query
    rows0, conf0, labels0 = jobs[0][:3]
    jobs.append((rows0[:2], conf0[:2], labels0[:2], out.masked_counts[0], out.viewed_counts, g.n))   # too few masks to aggregate
    single = [stage2.aggregate_and_filter(*j, cfg) for j in jobs]
    batch = stage2.aggregate_and_filter_batch(jobs, cfg)
    assert len(batch) == len(jobs)
    assert sum(r[1] for r in single) >= 2 and not single[-2][1]
    for k, (a, b) in enumerate(zip(single, batch)):
        assert a[1] == b[1], f"job {k}: reached_filter differs"
        _same(a[0], b[0], f"batch vs single, job {k}")


def test_batched_refinement_pass1_equals_one_call_per_query(cuda_dev):
    """bff_refine_match_batch against one bff_refine_match call per query (each on its own copy of the stage-1 rows),
    with a query whose stage 2 is empty."""
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops, refine, scene_io, stage2
    inp = bench.build_inputs("tiny3", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    cfg = inp["cfg"]
    out = g.lift(d, g.masks2d(d))
    s1 = g.stage1_rows(d)
    res2 = []
    for qi in range(g.nq):
        idx = g.row_index_dev[qi] if g.row_index_dev is not None else None
        r, _ = stage2.aggregate_and_filter(out.rows if idx is None else out.rows[idx], g.conf_q[qi], g.labels_q[qi], out.masked_counts[qi],
                                           out.viewed_counts, g.n, cfg, keep_rows=True)
        res2.append(r)
    res2.append(scene_io.empty_stage2())
    names = list(g.names) + [g.names[0]]
    # the device-decision pass 1 (what pass1_scene takes by default) ...
    on_device = [refine.pass1_scene(inp["stage1"], r, nm, cfg["stage1_iou_thres"], bench.fake_similarity, cuda_dev, s1_rows=s1.clone())
                 for r, nm in zip(res2, names)]
    # ... and the host-decision one, single and batched
    refine.DEVICE = False
    try:
        single = [refine.pass1_scene(inp["stage1"], r, nm, cfg["stage1_iou_thres"], bench.fake_similarity, cuda_dev, s1_rows=s1.clone())
                  for r, nm in zip(res2, names)]
    finally:
        refine.DEVICE = True
    for k, (a, b) in enumerate(zip(on_device, single)):
        assert a["sims"] == b["sims"] and torch.equal(torch.as_tensor(a["ious"]), torch.as_tensor(b["ious"])), k
        _same(refine.pass2_scene(a, names[k], cfg["refiment_iou_thres"], 0.0), refine.pass2_scene(b, names[k], cfg["refiment_iou_thres"], 0.0),
              f"device pass 1 + 2 vs host pass 1 + 2, query {k}")
    batch = refine.pass1_scene_batch(inp["stage1"], res2, names, cfg["stage1_iou_thres"], bench.fake_similarity, cuda_dev,
                                     [s1.clone() for _ in names])
    assert sum(len(p["sims"]) > 0 for p in single) >= 2
    for k, (a, b) in enumerate(zip(single, batch)):
        assert a["sims"] == b["sims"], k
        assert torch.equal(torch.as_tensor(a["ious"]), torch.as_tensor(b["ious"])), k
        assert torch.equal(a["other"], b["other"]), k
        if a["m1"] is not None:
            assert torch.equal(a["m1"], b["m1"]) and torch.equal(a["m2"], b["m2"]), k
            assert torch.equal(torch.as_tensor(a["conf2"]), torch.as_tensor(b["conf2"])), k
        fa = refine.pass2_scene(a, names[k], cfg["refiment_iou_thres"], 0.0)
        fb = refine.pass2_scene(b, names[k], cfg["refiment_iou_thres"], 0.0)
        _same(fa, fb, f"pass 2 after batched pass 1, query {k}")
