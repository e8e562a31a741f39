"""CPU, gloo, world 2 and 3: the round-2 multi-GPU algorithm with the oracle port standing in for
the kernels -- frames interleaved over the ranks and lifted over ALL points of the NaN-padded cloud (dist.ShardPlan),
row slices handed to the owners of the points, votes all-reduced, then aggregation / point filter / overlap resolution /
row filter on every rank's POINT SLICE with only the additive integers (intersections, popcounts) all-reduced and every
decision taken redundantly; then both refinement passes on the same slices (the port's own code with its IoU function fed
by all-reduced integers). The saved stage-2 masks and the final masks, re-assembled from the ranks' column slices, must
equal the single-process results of the port on the same scene. (The CUDA form of the same algorithm is checked on hardware by
tests/test_sharded_gpu.py; this test pins the host-side arithmetic -- padding, slice bounds, frame interleave, scene row
order -- and the claim that point-sharding changes no decision.)"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _paths():
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.join(here, ".."), os.path.join(here, "..", "oracle"), os.path.join(here, "..", "beyond-fixed-forms_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)


def _scene(occurrence):
    _paths()
    import bench
    inp = bench.build_inputs("tiny", 0, 1, torch.device("cpu"), seed=2000)
    cfg = dict(inp["cfg"])
    if occurrence:
        cfg.update(if_occurance_threshold=True, if_detected_ratio_threshold=False)
    per_query, view_frames = bench.cpu_frames_of(inp, 10 ** 9)
    have = {r["frame_id"][:-4] for r in inp["records"]}
    mask_frame_index = [i for i, fid in enumerate(inp["frame_ids"]) if fid in have]
    assert len(mask_frame_index) == len(per_query[0])
    return inp, cfg, per_query[0], mask_frame_index, view_frames


def _allreduce(t):
    import torch.distributed as dist
    t = t.to(torch.int64).contiguous()
    dist.all_reduce(t)
    return t


def _sharded_stage2(rank, world, occurrence):
    """What one rank computes; returns (its column slice of the saved rows, conf, classes, plan, diagnostics)."""
    import torch.distributed as dist
    import reference_port as RP
    from bff_b200.dist import ShardPlan
    inp, cfg, mask_frames, mask_frame_index, view_frames = _scene(occurrence)
    n = inp["xyz"].shape[0]
    plan = ShardPlan(n, world, rank)
    x4 = np.zeros((n, 4), np.float32)
    x4[:, :3] = inp["xyz"].astype(np.float32)
    padded = plan.padded_cloud(x4)
    assert padded.shape == (plan.n_padded, 4) and plan.n_padded % (32 * world) == 0 and plan.n_padded >= n
    assert np.isnan(padded[n:, :3]).all() and np.array_equal(padded[:n], x4)
    xyz_pad = padded[:, :3].astype(np.float64)
    # 1. this rank's frames: r, r + G, ... of the whole trajectory (mask frames are the ones among them that carry masks)
    mine = [(i, fr) for i, fr in zip(mask_frame_index, mask_frames) if i % world == rank]
    views = [fr for i, fr in enumerate(view_frames) if i % world == rank]
    with np.errstate(all="ignore"):
        rows_r, mc_r, vc_r = RP.lift_scene(xyz_pad, inp["K"], [fr for _, fr in mine], views)
    assert not bool(rows_r[:, n:].any()) and float(mc_r[n:].sum()) == 0 and float(vc_r[n:].sum()) == 0     # NaN points: never visible
    # 2. votes: order-free integer sums
    votes = _allreduce(torch.stack([mc_r, vc_r]))
    mc, vc = votes[0].float(), votes[1].float()
    # 3. the exchange: every rank ends up with ALL rows restricted to ITS points, in scene (frame) order
    meta = [(i, k, float(fr["conf"][k]), fr["labels"][k]) for i, fr in mine for k in range(len(fr["labels"]))]
    box = [None] * world
    dist.all_gather_object(box, (np.packbits(rows_r.numpy(), axis=1), meta))
    p0, p1 = plan.point0, plan.point0 + plan.slice_points
    blocks, metas = [], []
    for packed, m in box:
        full = np.unpackbits(packed, axis=1)[:, :plan.n_padded].astype(bool) if len(m) else np.zeros((0, plan.n_padded), bool)
        blocks.append(full[:, p0:p1])
        metas += m
    order = sorted(range(len(metas)), key=lambda j: (metas[j][0], metas[j][1]))
    ins = torch.from_numpy(np.concatenate(blocks)[order])
    conf = torch.tensor([metas[j][2] for j in order])
    labels = [metas[j][3] for j in order]
    # 4. aggregation on the slice: intersections and popcounts are summed over the ranks, the decisions are local
    a = ins.float()
    inter = _allreduce(a @ a.T).float()
    pop = _allreduce(a.sum(1)).float()
    iou = inter / (pop.unsqueeze(1) + pop.unsqueeze(0) - inter)
    merge = RP.same_label(labels) & (iou > cfg["iou_thres"])
    groups = [g for g in RP.components(merge.float()) if len(g) >= cfg["min_aggragated_masks"] and g != []]
    if not groups:
        return None, None, [], plan, dict(n_rows=len(labels), n_groups=0)
    agg = torch.stack([ins[g].any(0) for g in groups])
    conf_a = torch.tensor([sum(conf[k] for k in g) / len(g) for g in groups])
    labels_a = [labels[g[0]] for g in groups]
    keep = RP.point_filter(mc.clone(), vc, cfg)[p0:p1]                 # the threshold needs the votes of ALL points
    before = _allreduce(agg.sum(1))
    overlap = _allreduce((agg.float() @ agg.float().T))               # solve_overlapping's pair list: any shared point anywhere
    size = [len(g) for g in groups]
    for i in range(len(groups)):
        for j in range(i + 1, len(groups)):
            if overlap[i, j] > 0:
                if size[i] > size[j]:
                    agg[j] &= ~agg[i]
                else:
                    agg[i] &= ~agg[j]
    agg &= keep.unsqueeze(0)
    after = _allreduce(agg.sum(1))
    sel = (after > cfg["remove_small_masks"]) & (after > cfg["remove_filtered_masks"] * before)
    return agg[sel], conf_a[sel], [l for l, s in zip(labels_a, sel.tolist()) if s], plan, dict(n_rows=len(labels), n_groups=len(groups))


def _sharded_refine(inp, cfg, plan, s2_local):
    """Both refinement passes on this rank's point slice: the port's own code, with the one function that needs ALL
    points -- the IoU between two mask sets -- fed by all-reduced intersections and popcounts."""
    import bench
    import reference_port as RP
    from bff_b200.refine import STAGE1_CLASS_NAMES
    p0, p1 = plan.point0, plan.point0 + plan.slice_points
    s1 = np.zeros((len(inp["stage1"]["ins"]), plan.n_padded), np.uint8)
    s1[:, :plan.n_points] = np.stack([RP.rle_to_mask(r) for r in inp["stage1"]["ins"]])
    s1 = torch.from_numpy(np.ascontiguousarray(s1[:, p0:p1]))
    labels = [STAGE1_CLASS_NAMES[int(i)] for i in inp["stage1"]["final_class"]]

    def iou_all_points(m1, m2):
        a, b = m1.float(), m2.float()
        inter = _allreduce(a @ b.T).float()
        sa, sb = _allreduce(a.sum(dim=-1, keepdim=True)).float(), _allreduce(b.sum(dim=-1, keepdim=True)).float()
        return (inter / (sa + sb.T - inter)).T

    whole, RP.iou_between = RP.iou_between, iou_all_points
    try:
        query = inp["queries"][0]
        pass1 = RP.refine_pass1(s1, labels, s2_local, query, cfg["stage1_iou_thres"], bench.fake_similarity)
        thr = RP.sim_threshold([pass1["sims"]], cfg["refinment_sim_percentile"]) if len(pass1["sims"]) else 0.0
        return RP.refine_pass2(pass1, query, cfg["refiment_iou_thres"], thr)
    finally:
        RP.iou_between = whole


def _gather_columns(local, plan):
    import torch.distributed as dist
    box = [None] * plan.world
    dist.all_gather_object(box, None if local is None else np.asarray(local))
    return None if local is None else torch.from_numpy(np.concatenate(box, axis=1)[:, :plan.n_points])


def _worker(rank, world, port, out_dir, occurrence):
    _paths()
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    ins, conf, classes, plan, info = _sharded_stage2(rank, world, occurrence)
    full = _gather_columns(None if ins is None else ins.numpy(), plan)
    final = None
    if ins is not None:
        inp, cfg, *_ = _scene(occurrence)
        out = _sharded_refine(inp, cfg, plan, {"ins": ins, "conf": conf, "final_class": classes})
        final = {"ins": _gather_columns(out["ins"].numpy(), plan) if len(out["ins"]) else [], "conf": out["conf"], "final_class": out["final_class"]}
    if rank == 0:
        torch.save({"ins": full, "conf": conf, "final_class": classes, "info": info, "final": final}, os.path.join(out_dir, "sharded.pth"))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,occurrence", [(2, False), (3, False), (2, True)])
def test_point_sharded_stage2_and_refinement_equal_single_process(tmp_path, world, occurrence):
    _paths()
    import bench
    import reference_port as RP
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), occurrence), nprocs=world, join=True)
    got = torch.load(os.path.join(str(tmp_path), "sharded.pth"), weights_only=False)
    inp, cfg, mask_frames, _, view_frames = _scene(occurrence)
    want = RP.stage2_scene(inp["xyz"], inp["K"], mask_frames, view_frames, cfg)
    assert got["info"]["n_rows"] > 8 and got["info"]["n_groups"] >= 1
    assert want["ins"].shape[0] >= 1, "the scene must produce masks for the comparison to mean anything"
    assert got["final_class"] == want["final_class"]
    assert torch.equal(got["ins"], want["ins"])
    assert torch.equal(got["conf"], want["conf"])
    # ... and after both refinement passes (stage-1 matching, merges, similarity threshold, selection)
    inp["cfg"] = cfg
    want_final = bench.cpu_step(inp, [mask_frames], view_frames)[0]
    final = got["final"]
    assert len(want_final["final_class"]) >= 1 and final["final_class"] == want_final["final_class"]
    assert torch.equal(final["ins"], want_final["ins"])
    assert torch.equal(torch.as_tensor(final["conf"]), torch.as_tensor(want_final["conf"]))


def test_shard_plan_slices_tile_the_padded_cloud():
    _paths()
    from bff_b200.dist import ShardPlan
    for n in (1, 31, 32, 33, 20_000, 250_000, 2_000_003):
        for world in (1, 2, 3, 8):
            plans = [ShardPlan(n, world, r) for r in range(world)]
            p = plans[0]
            assert p.words == (n + 31) // 32 and p.slice_words * world >= p.words and (p.slice_words - 1) * world < max(p.words, world)
            assert p.n_padded == world * p.slice_points and p.slice_points == 32 * p.slice_words
            assert [q.point0 for q in plans] == [r * p.slice_points for r in range(world)]
            assert plans[-1].point0 + p.slice_points == p.n_padded >= n
