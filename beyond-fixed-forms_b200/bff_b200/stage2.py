"""Stage 2 driver: lifting + aggregation + filtering of one scene / one class, the body of
tools/projection_2d_to_3d.py __main__ (:336-634) on the GPU path."""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _log, aggregate, lift as lift_mod, scene_io, schedule
from ._log import log


FUSED = True     # bff_stage2_aggregate (C++ decisions next to the launches); False = one Python/torch call per step
MAX_DEVICE_ROWS, MAX_FUSED_ROWS = 25000, 46340   # lifted-mask counts bff_stage2_device / bff_stage2_aggregate accept (include/bff.h); beyond: next path down
DEVICE = True    # bff_stage2_device (every decision a kernel, no host round trip inside the call); takes precedence over FUSED


def cfg_get(cfg, key):
    return cfg[key] if isinstance(cfg, dict) else getattr(cfg, key)


def needs_viewed_counts(cfg) -> bool:
    """viewed_counts are only computed on the detected-ratio branch (projection_2d_to_3d.py:512,524)."""
    return (not cfg_get(cfg, "if_occurance_threshold")) and bool(cfg_get(cfg, "if_detected_ratio_threshold"))


_CFG_KEYS = ("iou_thres", "min_aggragated_masks", "if_occurance_threshold", "occurance_threshold",
             "if_detected_ratio_threshold", "detected_ratio_threshold", "remove_filtered_masks", "remove_small_masks")


def _finish_fused(fused, n_points: int, c, keep_rows: bool, to_host: bool):
    """Row filter + saved dict for one bff_stage2_aggregate result (projection_2d_to_3d.py:592-634)."""
    if fused is None:
        log("No 3d masks detected after aggregation")
        return scene_io.empty_stage2(), False
    agg, conf_agg, labels_agg, groups, before64, after64 = fused
    log("before filtering", torch.Size([agg.shape[0], n_points]))
    result, out_rows = aggregate.select_rows(agg, conf_agg, labels_agg, before64, after64, n_points, c, to_host)
    if keep_rows:
        result["rows"] = out_rows        # packed CUDA rows for an in-process refinement (never saved)
    return result, True


def aggregate_and_filter(rows: torch.Tensor, conf_list, labels, mc: torch.Tensor, vc: Optional[torch.Tensor],
                         n_points: int, cfg, keep_rows: bool = False, to_host: bool = True, pairs=None, n_frames: Optional[int] = None):
    """Everything after the frame loop (projection_2d_to_3d.py:463-634) for one query.
    Returns (dict to save, reached_filter) -- the reference only writes its scene checkpoint when
    the filter stage is reached (:580-581). `n_frames` (the frames that were lifted: an upper bound of the viewed
    counts) enables the device-resident path."""
    c = {k: cfg_get(cfg, k) for k in _CFG_KEYS}
    if len(conf_list) == 0:
        log("No 3d masks detected in backprojection!")
        return scene_io.empty_stage2(), False
    conf = conf_list if isinstance(conf_list, torch.Tensor) else torch.tensor(conf_list)    # :484-486
    if DEVICE and n_frames is not None and pairs is None and c["min_aggragated_masks"] >= 1 and conf.dtype == torch.float32 \
            and rows.shape[0] <= MAX_DEVICE_ROWS:
        from . import devpipe, ops
        s2 = devpipe.stage2_device(rows, ops.small_to_device(conf, rows.device), labels, mc, vc, n_points, c,
                                   mc_cap=rows.shape[0], vc_cap=int(n_frames))
        return s2.saved_dict(to_host=to_host, keep_rows=keep_rows)
    if FUSED and c["min_aggragated_masks"] >= 1 and conf.dtype == torch.float32 and rows.shape[0] <= MAX_FUSED_ROWS:
        fused = aggregate.aggregate_filter_fused(rows, conf, labels, mc, vc, n_points, c, pairs=pairs)
        return _finish_fused(fused, n_points, c, keep_rows, to_host)
    agg, conf_agg, labels_agg, groups = aggregate.aggregate_rows(rows, conf, labels, c["iou_thres"], c["min_aggragated_masks"])
    if agg is None:
        log("No 3d masks detected after aggregation")
        return scene_io.empty_stage2(), False
    keep = aggregate.keep_points(mc, vc, c)
    log("before filtering", torch.Size([agg.shape[0], n_points]))
    result, out_rows = aggregate.filter_and_select(agg, conf_agg, labels_agg, groups, keep, n_points, c, to_host)
    if keep_rows:
        result["rows"] = out_rows
    return result, True


def aggregate_and_filter_batch(jobs, cfg, keep_rows: bool = False, to_host: bool = True):
    """aggregate_and_filter for several queries of one scene: jobs = [(rows, conf_list, labels, mc, vc, n_points), ...].
    The fused jobs go through ONE bff_stage2_aggregate_batch call (shared synchronisations); a job the fused path does
    not take (nothing lifted, non-float32 confidences, min_aggragated_masks < 1) is handled on its own.
    Returns [(dict to save, reached_filter), ...] in job order, equal to one aggregate_and_filter call per job."""
    c = {k: cfg_get(cfg, k) for k in _CFG_KEYS}
    out = [None] * len(jobs)
    fused_at, fused_jobs = [], []
    for k, (rows, conf_list, labels, mc, vc, n_points) in enumerate(jobs):
        conf = conf_list if isinstance(conf_list, torch.Tensor) else (torch.tensor(conf_list) if len(conf_list) else None)
        if conf is not None and len(conf_list) and FUSED and c["min_aggragated_masks"] >= 1 and conf.dtype == torch.float32:
            fused_at.append(k)
            fused_jobs.append((rows, conf, labels, mc, vc, n_points))
        else:
            out[k] = aggregate_and_filter(rows, conf_list, labels, mc, vc, n_points, cfg, keep_rows, to_host)
    for k, fused in zip(fused_at, aggregate.aggregate_filter_fused_batch(fused_jobs, c)):
        out[k] = _finish_fused(fused, jobs[k][5], c, keep_rows, to_host)
    return out


def load_scene(cfg, cls: str, scene_id: str, device, with_frames: bool = True) -> dict:
    """Everything process_scene reads from disk for one (class, scene): intrinsics, cloud, 2D-mask records, the frame
    plan and (single-GPU path) the frames themselves. Pure host work apart from the cached cv2.resize gate, so the
    scene loop runs it for scene i+1 on a background thread while scene i is on the GPU."""
    torch.cuda.set_device(device)
    width, height = int(cfg_get(cfg, "width_2d")), int(cfg_get(cfg, "height_2d"))
    scene_2d_dir = cfg_get(cfg, "scene_2d_dir")
    intr = scene_io.load_intrinsics(scene_2d_dir, scene_id)
    xyz = scene_io.load_cloud(cfg_get(cfg, "scene_npy_dir"), scene_id)
    records = scene_io.torch_load(os.path.join(cfg_get(cfg, "mask_2d_dir"), cls, f"{scene_id}.pth"))
    q = lift_mod.QueryMasks.from_file_records(cls, records)
    want_viewed = needs_viewed_counts(cfg)
    conf_list, labels = q.flat_conf_and_labels()
    # the reference skips the viewed-count pass when nothing was lifted or nothing aggregates; its
    # result is unused in those cases, so computing it in the same launch changes nothing on disk
    view_ids = scene_io.list_view_frames(scene_2d_dir, scene_id, int(cfg_get(cfg, "downsample_ratio"))) if (want_viewed and conf_list) else None
    frame_ids, viewed, per_frame = lift_mod.plan_frames([q], view_ids)
    frames = scene_io.load_frames(scene_2d_dir, scene_id, frame_ids, viewed, width, height, device) if with_frames else None
    return dict(intr=intr, xyz=xyz, q=q, conf_list=conf_list, labels=labels, want_viewed=want_viewed, frame_ids=frame_ids,
                viewed=viewed, per_frame=per_frame, frames=frames, width=width, height=height)


def process_scene(cfg, cls: str, scene_id: str, device, dist_ctx=None, loaded: Optional[dict] = None):
    """One iteration of the scene loop (projection_2d_to_3d.py:365-634).
    Returns (dict to save, reached_filter). `loaded`: the result of load_scene when it was prefetched."""
    from . import dist as dist_mod
    log("Working on", scene_id, "class", cls)
    sharded = dist_ctx is not None and dist_ctx.world_size > 1
    L = loaded if loaded is not None else load_scene(cfg, cls, scene_id, device, with_frames=not sharded)
    xyz, q, conf_list, labels, want_viewed = L["xyz"], L["q"], L["conf_list"], L["labels"], L["want_viewed"]
    if sharded:
        lifted = dist_mod.lift_sharded(dist_ctx, xyz, L["intr"], cfg_get(cfg, "scene_2d_dir"), scene_id, L["frame_ids"], L["viewed"],
                                       L["per_frame"], [q], L["height"], L["width"], want_viewed and bool(conf_list), device)
    else:
        lifted = lift_mod.lift(scene_io.cloud_for_device(xyz, device), L["intr"], L["frames"], L["per_frame"], [q], L["height"],
                               L["width"], want_viewed and bool(conf_list))
    pairs = None
    if sharded and lifted.rows[0].shape[0] >= dist_mod.PAIRS_MIN_ROWS:
        pairs = dist_mod.sharded_pairs(dist_ctx, lifted.rows[0])      # the quadratic IoU step, one row block per rank
    return aggregate_and_filter(lifted.rows[0], conf_list, labels, lifted.masked_counts[0], lifted.viewed_counts,
                                xyz.shape[0], cfg, pairs=pairs, n_frames=lifted.n_frames)


def scene_loop(cfg, cls: str, scene_ids, device):
    """Single-GPU scene loop with the next scene's files loaded on a background thread (PNG decode, text parsing and
    torch.load release the GIL most of the time). Yields (scene_id, result, reached_filter) in order."""
    from concurrent.futures import ThreadPoolExecutor
    scene_ids = list(scene_ids)
    if not scene_ids:
        return
    with ThreadPoolExecutor(max_workers=1, thread_name_prefix="bff-ingest") as ex:
        nxt = ex.submit(load_scene, cfg, cls, scene_ids[0], device)
        for i, scene_id in enumerate(scene_ids):
            loaded = nxt.result()
            if i + 1 < len(scene_ids):
                nxt = ex.submit(load_scene, cfg, cls, scene_ids[i + 1], device)
            result, reached_filter = process_scene(cfg, cls, scene_id, device, None, loaded)
            yield scene_id, result, reached_filter


def process_scene_queries(cfg, classes, scene_id: str, device):
    """Several text queries of ONE scene in one pass (BASELINE config 3): projection,
    visibility and viewed_counts are computed once and shared, the 2D masks of all queries are gathered in
    the same launch (per-query vote buffers), then each query is aggregated and filtered on its own rows.
    Returns {cls: (dict to save, reached_filter)}; every dict equals what process_scene gives for that class."""
    width, height = int(cfg_get(cfg, "width_2d")), int(cfg_get(cfg, "height_2d"))
    scene_2d_dir = cfg_get(cfg, "scene_2d_dir")
    intr = scene_io.load_intrinsics(scene_2d_dir, scene_id)
    xyz = scene_io.load_cloud(cfg_get(cfg, "scene_npy_dir"), scene_id)
    queries = []
    for cls in classes:
        records = scene_io.torch_load(os.path.join(cfg_get(cfg, "mask_2d_dir"), cls, f"{scene_id}.pth"))
        queries.append(lift_mod.QueryMasks.from_file_records(cls, records))
    want_viewed = needs_viewed_counts(cfg)
    any_masks = any(len(r) for q in queries for r in q.rles)
    view_ids = scene_io.list_view_frames(scene_2d_dir, scene_id, int(cfg_get(cfg, "downsample_ratio"))) if (want_viewed and any_masks) else None
    frame_ids, viewed, per_frame = lift_mod.plan_frames(queries, view_ids)
    out = {}
    if not frame_ids:
        return {q.query: aggregate_and_filter(None, [], [], None, None, xyz.shape[0], cfg) for q in queries}
    frames = scene_io.load_frames(scene_2d_dir, scene_id, frame_ids, viewed, width, height, device)
    lifted = lift_mod.lift(scene_io.cloud_for_device(xyz, device), intr, frames, per_frame, queries, height, width,
                           want_viewed and any_masks)
    def job(qi):
        conf_list, labels = queries[qi].flat_conf_and_labels()
        return (lifted.rows[qi], conf_list, labels, lifted.masked_counts[qi], lifted.viewed_counts, xyz.shape[0])
    if DEVICE:
        # every query's chain is enqueue-only (bff_stage2_device); the results are read back once they are all queued
        results = [aggregate_and_filter(*job(qi), cfg, n_frames=lifted.n_frames) for qi in range(len(queries))]
    else:
        # all queries advance phase by phase inside one call and share its synchronisations
        results = aggregate_and_filter_batch([job(qi) for qi in range(len(queries))], cfg)
    for q, r in zip(queries, results):
        out[q.query] = r
    return out


def main_multi(argv=None):
    """tools/projection_2d_to_3d_multi.py: `--config C --cls q1 q2 ...` -- the same outputs and checkpoints as one
    reference run per class, with the per-scene projection shared by all classes."""
    import argparse
    import yaml
    parser = argparse.ArgumentParser(description="BeyondFF lifting for several classes at once")
    parser.add_argument("--config", type=str, required=True)
    parser.add_argument("--cls", type=str, nargs="+", required=True)
    args = parser.parse_args(argv)
    _log.VERBOSE = True
    cfg = yaml.safe_load(open(args.config, "r").read())
    if not torch.cuda.is_available():
        raise RuntimeError("bff_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    device = torch.device("cuda", 0)
    scenes = None
    for cls in args.cls:                                   # the classes must list the same scenes
        names = sorted(s for s in os.listdir(os.path.join(cfg["mask_2d_dir"], cls)) if s.endswith("_00.pth"))
        if scenes is not None and names != scenes:
            raise ValueError(f"class {cls!r} lists other scenes than {args.cls[0]!r}")
        scenes = names
    ckpts = {cls: scene_io.read_checkpoint(f"checkpoints/projection_2d_to_3d_checkpoint_{cls}.yaml") for cls in args.cls}
    for s in scenes:
        scene_id = s[:-4]
        for cls, (result, reached_filter) in process_scene_queries(cfg, args.cls, scene_id, device).items():
            scene_io.save_result(result, cfg["mask_3d_dir"], cls, scene_id)
            if reached_filter:
                ckpts[cls][scene_id] = True
                scene_io.write_checkpoint(f"checkpoints/projection_2d_to_3d_checkpoint_{cls}.yaml", ckpts[cls])
    return 0


def main(argv=None):
    """CLI body of tools/projection_2d_to_3d.py: same arguments, config keys, outputs, checkpoints."""
    import argparse
    import yaml
    parser = argparse.ArgumentParser(description="Configuration Open3DIS")
    parser.add_argument("--config", type=str, required=True, help="Config")
    parser.add_argument("--cls", type=str, required=True, help="Class")
    args = parser.parse_args(argv)
    _log.VERBOSE = True
    cfg = yaml.safe_load(open(args.config, "r").read())
    if not torch.cuda.is_available():
        raise RuntimeError("bff_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    from . import dist as dist_mod
    dist_ctx = dist_mod.init_from_env()
    device = torch.device("cuda", dist_ctx.local_rank if dist_ctx else 0)
    torch.cuda.set_device(device)
    if dist_ctx is not None:
        dist_mod.bind_to_gpu_numa(dist_ctx.local_rank)
    cls = args.cls
    ckpt_path = f"checkpoints/projection_2d_to_3d_checkpoint_{cls}.yaml"     # :320-322
    ckpt = scene_io.read_checkpoint(ckpt_path)
    seg_dir = os.path.join(cfg["mask_2d_dir"], cls)
    scenes = sorted(s for s in os.listdir(seg_dir) if s.endswith("_00.pth"))  # :361-363
    if dist_ctx is not None and dist_ctx.world_size > 1 and schedule.dist_mode(len(scenes), dist_ctx.world_size) == "scene":
        # whole scenes per rank (longest first onto the least loaded rank), no data-path collective; rank 0 merges the
        # checkpoint entries at the end
        import torch.distributed as tdist
        mine = schedule.lpt_assign([schedule.scene_cost(cfg, s[:-4]) for s in scenes], dist_ctx.world_size)[dist_ctx.rank]
        reached = []
        for scene_id, result, reached_filter in scene_loop(cfg, cls, [scenes[i][:-4] for i in mine], device):
            scene_io.save_result(result, cfg["mask_3d_dir"], cls, scene_id)
            if reached_filter:
                reached.append(scene_id)
        everyone = [None] * dist_ctx.world_size
        tdist.all_gather_object(everyone, reached)
        if dist_ctx.rank == 0:
            for scene_id in sorted(x for part in everyone for x in part):
                ckpt[scene_id] = True
            scene_io.write_checkpoint(ckpt_path, ckpt)
        dist_ctx.finalize()
        return 0
    if dist_ctx is None or dist_ctx.world_size == 1:
        for scene_id, result, reached_filter in scene_loop(cfg, cls, [s[:-4] for s in scenes], device):
            scene_io.save_result(result, cfg["mask_3d_dir"], cls, scene_id)
            if reached_filter:
                ckpt[scene_id] = True
                scene_io.write_checkpoint(ckpt_path, ckpt)
        if dist_ctx is not None:
            dist_ctx.finalize()
        return 0
    for s in scenes:
        scene_id = s[:-4]
        result, reached_filter = process_scene(cfg, cls, scene_id, device, dist_ctx)
        if dist_ctx is None or dist_ctx.rank == 0:
            scene_io.save_result(result, cfg["mask_3d_dir"], cls, scene_id)
            if reached_filter:
                ckpt[scene_id] = True
                scene_io.write_checkpoint(ckpt_path, ckpt)
    if dist_ctx is not None:
        dist_ctx.finalize()
    return 0
