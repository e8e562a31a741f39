"""Frame-sharded multi-GPU lifting (one process per GPU, torch.distributed over NCCL/NVLink).

The reference has no distributed path. Frames are independent until votes are
summed and lifted rows are collected, so rank r lifts a contiguous block of the planned frames
and exactly one exchange step follows:
  * all_reduce(SUM) over int32 [viewed_counts | masked_counts(Q,N)]  -- integer, order-free, bit-exact
  * all_gather of the packed lifted rows in rank order (= launch order of the 1-GPU path)
Aggregation then runs on every rank (tiny host work + a few kernels). One GPU takes no collective.
"""
from __future__ import annotations

import dataclasses
import os
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from . import lift as lift_mod, scene_io


@dataclasses.dataclass
class DistCtx:
    rank: int
    world_size: int
    local_rank: int
    owns_group: bool = False

    def finalize(self):
        import torch.distributed as dist
        if self.owns_group and dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()


def init_from_env(backend: Optional[str] = None) -> Optional[DistCtx]:
    """torchrun-style environment (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*). None for 1 process."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return None
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    owns = False
    if not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
        owns = True
    return DistCtx(rank, world, local, owns)


def bind_to_gpu_numa(device_index: int) -> Optional[list]:
    """Pin this process (and thereby the first-touch placement of the pinned host buffers it allocates next) to the CPU
    cores NVML reports as local to GPU `device_index`. With one rank per GPU on a multi-socket host, staging buffers on
    the far socket make every upload cross the inter-socket link, which caps the aggregate host->device bandwidth of an
    8-GPU box well below 8 x PCIe. Returns the CPU list, or None when NVML / affinity control is unavailable
    (BFF_NO_NUMA_BIND=1 disables it)."""
    if os.environ.get("BFF_NO_NUMA_BIND") == "1" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(visible.split(",")[device_index]) if visible and all(v.strip().isdigit() for v in visible.split(",")) else device_index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = [w * 64 + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


def shard_bounds(n_items: int, world_size: int, rank: int):
    """Contiguous block [lo, hi) of rank `rank`: the first (n mod world) ranks get one extra item."""
    base, extra = divmod(n_items, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def exchange(ctx: DistCtx, rows_local: torch.Tensor, rows_per_rank: Sequence[int], mc: torch.Tensor,
             vc: Optional[torch.Tensor]):
    """The one exchange step. rows_local (k_r, Wd) int32; counts int32. Returns (all rows in rank
    order, summed mc, summed vc)."""
    import torch.distributed as dist
    n = mc.shape[-1]
    packed = torch.cat([mc.reshape(-1), vc.reshape(-1)]) if vc is not None else mc.reshape(-1).clone()
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    mc_sum = packed[: mc.numel()].reshape(mc.shape)
    vc_sum = packed[mc.numel():].reshape(n) if vc is not None else None
    kmax = max(rows_per_rank) if len(rows_per_rank) else 0
    wd = rows_local.shape[1]
    if kmax == 0:
        return rows_local[:0], mc_sum, vc_sum
    pad = torch.zeros((kmax, wd), dtype=rows_local.dtype, device=rows_local.device)
    pad[: rows_local.shape[0]] = rows_local
    gathered = [torch.empty_like(pad) for _ in range(ctx.world_size)]
    dist.all_gather(gathered, pad)
    rows = torch.cat([gathered[r][: rows_per_rank[r]] for r in range(ctx.world_size)], dim=0)
    return rows, mc_sum, vc_sum


def lift_sharded(ctx: DistCtx, xyz: np.ndarray, intrinsics: np.ndarray, scene_2d_dir: str, scene_id: str,
                 frame_ids: List[str], viewed: List[bool], per_frame: List[list],
                 queries: Sequence[lift_mod.QueryMasks], height: int, width: int, want_viewed: bool, device,
                 lift_fn: Optional[Callable] = None, load_fn: Optional[Callable] = None) -> lift_mod.LiftedScene:
    """Lift this rank's contiguous block of the planned frames, then exchange.
    `lift_fn` / `load_fn` default to the CUDA kernel and the on-disk loader; the gloo CPU tests
    inject the oracle there to check the sharding and exchange logic without a GPU."""
    f = len(frame_ids)
    lo, hi = shard_bounds(f, ctx.world_size, ctx.rank)
    load_fn = load_fn or (lambda ids, vw: scene_io.load_frames(scene_2d_dir, scene_id, ids, vw, width, height, device))
    lift_fn = lift_fn or lift_mod.lift_frames
    frames = load_fn(frame_ids[lo:hi], viewed[lo:hi])
    rles, mask_query, offsets = lift_mod.mask_layout(per_frame, queries, lo, hi)
    xyz_dev = scene_io.cloud_for_device(xyz, device)
    out = lift_fn(xyz_dev, intrinsics, frames, rles, mask_query, offsets, len(queries), height, width, want_viewed)
    rows_per_rank = []
    for r in range(ctx.world_size):
        a, b = shard_bounds(f, ctx.world_size, r)
        rows_per_rank.append(sum(len(queries[qi].rles[fi]) for s in range(a, b) for (qi, fi) in per_frame[s]))
    rows, mc, vc = exchange(ctx, out.rows, rows_per_rank, out.masked_counts, out.viewed_counts if want_viewed else None)
    return lift_mod.LiftedScene(xyz.shape[0], lift_mod.split_rows(rows, per_frame, queries), mc, vc, f)


PAIRS_MIN_ROWS = int(os.environ.get("BFF_PAIRS_MIN_ROWS", "512"))   # below this the one-GPU call is cheaper than two more collectives


def sharded_pairs(ctx: DistCtx, rows: torch.Tensor, iou_fn: Optional[Callable] = None):
    """All-pairs intersections of the gathered rows with the quadratic work split over the ranks: rank r computes one
    trapezoid of the upper triangle (bff_mask_iou), the int32 blocks are all-gathered in rank order. Integer results,
    so the assembled (n,n) matrix is bit-identical to the one-GPU call. Returns (inter (n,n) int32, pop (n,) int32)
    for aggregate_filter_fused(pairs=...). `iou_fn(block, rows) -> (inter, pop_block)` defaults to the CUDA kernel; the
    gloo CPU tests inject the oracle. (calculate_iou, projection_2d_to_3d.py:149-166, is O(Ins^2 N): at c4 it is
    the step that would otherwise be repeated on every rank.)"""
    import torch.distributed as dist
    from . import ops
    n = rows.shape[0]
    world = ctx.world_size
    # The matrix is symmetric: rank r computes rows [lo_r, hi_r) against columns [lo_r, n) only, and the boundaries
    # balance the AREA of those trapezoids (lo_r = n (1 - sqrt(1 - r/world))), so every rank does 1/world of the
    # upper triangle; the lower triangle is mirrored after the gather.
    bounds = [min(n, int(round(n * (1.0 - (1.0 - r / world) ** 0.5)))) for r in range(world)] + [n]
    lo, hi = bounds[ctx.rank], bounds[ctx.rank + 1]
    base = max(bounds[r + 1] - bounds[r] for r in range(world))         # all_gather needs one shape: pad to the tallest block
    block = torch.zeros((base, n), dtype=torch.int32, device=rows.device)
    pop_b = torch.zeros((base,), dtype=torch.int32, device=rows.device)
    if hi > lo:
        if iou_fn is None:
            inter, pa, _, _ = ops.mask_iou(rows[lo:hi], rows[lo:], want_iou=False)
        else:
            inter, pa = iou_fn(rows[lo:hi], rows[lo:])
        block[: hi - lo, lo:] = inter
        pop_b[: hi - lo] = pa
    inter_all = torch.empty((world * base, n), dtype=torch.int32, device=rows.device)
    pop_all = torch.empty((world * base,), dtype=torch.int32, device=rows.device)
    dist.all_gather_into_tensor(inter_all, block)
    dist.all_gather_into_tensor(pop_all, pop_b)
    keep = torch.cat([torch.arange(r * base, r * base + bounds[r + 1] - bounds[r], device=rows.device) for r in range(world)])
    upper = inter_all[keep]                                              # (n, n): entry (i, j) valid for j >= lo(block of i)
    row_lo = torch.cat([torch.full((bounds[r + 1] - bounds[r],), bounds[r], dtype=torch.long, device=rows.device) for r in range(world)])
    valid = torch.arange(n, device=rows.device)[None, :] >= row_lo[:, None]
    return torch.where(valid, upper, upper.T).contiguous(), pop_all[keep].contiguous()
