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
