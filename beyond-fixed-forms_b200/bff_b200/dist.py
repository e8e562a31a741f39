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


# ----------------------------------------------------------------------------- frame-sharded lift -> point-sharded pipelines
# Frames are independent only up to the votes and the lifted rows; everything after the frame loop (the all-pairs IoU,
# merges, filters, refinement) is element-wise over POINTS plus tiny decisions. So the work is handed over once: rank r
# lifts its block of frames over ALL points, writing its rows in G column slices (ops.lift_scene(row_slices=...)); one
# grouped send/recv gives every rank ALL rows of the scene restricted to ITS contiguous slice of the points; the votes are
# all-reduced. From there every rank runs the device-resident pipelines (devpipe.py) on its slice, the additive integer
# results (intersections, popcounts) are all-reduced inside those calls, and every rank takes the same decisions. No
# rank ever holds or computes the whole quadratic step. The cloud is padded with NaN points (never visible) to
# G * slice_words * 32 points so that all slices have the same width.
@dataclasses.dataclass
class ShardPlan:
    n_points: int
    world: int
    rank: int

    @property
    def words(self) -> int:
        return (self.n_points + 31) // 32

    @property
    def slice_words(self) -> int:
        return max(1, -(-self.words // self.world))

    @property
    def n_padded(self) -> int:
        return self.world * self.slice_words * 32

    @property
    def slice_points(self) -> int:
        return self.slice_words * 32

    @property
    def point0(self) -> int:
        return self.rank * self.slice_points

    def padded_cloud(self, xyz4: np.ndarray) -> np.ndarray:
        """(N,4) float32 -> (n_padded,4) with NaN points behind the cloud (a NaN point is never visible)."""
        out = np.full((self.n_padded, 4), np.nan, np.float32)
        out[: xyz4.shape[0]] = xyz4
        out[:, 3] = 0.0
        return out


class Comm:
    """The library's own NCCL communicator (bff_comm): created from torch.distributed's process group, which carries
    the 128-byte unique id from rank 0 to the others."""

    def __init__(self, ctx: DistCtx, device_index: int):
        import ctypes
        import torch.distributed as dist
        from . import _cabi
        self.ctx = ctx
        ident = ctypes.create_string_buffer(128)
        if ctx.rank == 0:
            _cabi.check(_cabi.lib().bff_comm_unique_id(ident), "bff_comm_unique_id")
        box = [bytes(ident.raw)]
        dist.broadcast_object_list(box, src=0)
        self._id = ctypes.create_string_buffer(box[0], 128)
        h = ctypes.c_void_p()
        _cabi.check(_cabi.lib().bff_comm_create(int(device_index), ctx.rank, ctx.world_size, self._id, ctypes.byref(h)), "bff_comm_create")
        self.handle = h

    def allreduce_(self, t: torch.Tensor, op: int = 0) -> torch.Tensor:
        from . import _cabi
        if t.dtype != torch.int32 or not t.is_contiguous():
            raise TypeError("allreduce_ needs a contiguous int32 tensor")
        _cabi.check(_cabi.lib().bff_comm_allreduce_i32(self.handle, _cabi.ptr(t), t.numel(), op, _cabi.current_stream()), "bff_comm_allreduce_i32")
        return t

    def exchange_rows(self, rows_sliced: torch.Tensor, rows_per_rank: Sequence[int]) -> torch.Tensor:
        """(G, n_local, slice_words) -> (sum(rows_per_rank), slice_words): all rows of the scene, this rank's points."""
        from . import _cabi
        g, n_local, sw = rows_sliced.shape
        rpr = np.asarray(rows_per_rank, np.int32)
        out = torch.empty((int(rpr.sum()), sw), dtype=torch.int32, device=rows_sliced.device)
        _cabi.check(_cabi.lib().bff_comm_exchange_rows(self.handle, _cabi.ptr(rows_sliced), n_local, _cabi.ptr(rpr), sw, _cabi.ptr(out),
                                                       _cabi.current_stream()), "bff_comm_exchange_rows")
        return out

    def allgather_rows(self, rows: torch.Tensor) -> torch.Tensor:
        """(k, slice_words) on every rank -> (G, k, slice_words) on every rank."""
        from . import _cabi
        rows = rows.contiguous()
        out = torch.empty((self.ctx.world_size,) + tuple(rows.shape), dtype=rows.dtype, device=rows.device)
        _cabi.check(_cabi.lib().bff_comm_allgather_bytes(self.handle, _cabi.ptr(rows), _cabi.ptr(out), rows.numel() * rows.element_size(),
                                                         _cabi.current_stream()), "bff_comm_allgather_bytes")
        return out

    def close(self):
        from . import _cabi
        if self.handle:
            _cabi.lib().bff_comm_destroy(self.handle)
            self.handle = None


class _RawCuda:
    """Minimal __cuda_array_interface__ holder: lets torch view device memory this package allocated itself."""

    def __init__(self, addr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(addr), False), "version": 3}


class PeerRows:
    """Receive buffers of the FUSED exchange: every rank owns one (n_total_rows, slice_words) int32 buffer allocated
    by libbff (cudaMalloc) and mapped into all other ranks with CUDA IPC, so that the lifting kernel of rank r stores
    slice p of its rows straight into rank p's buffer over NVLink (ops.lift_scene(peer_rows=...)): no separate row
    exchange, no send buffer. Protocol per scene: the owner zeroes its buffer, the ranks synchronise (a one-word
    all-reduce on the stream), every rank launches its lifting kernel, and the vote all-reduce that follows is also the
    barrier after which every rank's stores have landed. Raises BffError when IPC is unavailable (the caller then keeps
    the NCCL send/recv exchange)."""

    def __init__(self, ctx: DistCtx, n_total_rows: int, slice_words: int, device):
        import ctypes
        import torch.distributed as dist
        from . import _cabi
        self.ctx, self.shape = ctx, (int(n_total_rows), int(slice_words))
        nbytes = max(1, self.shape[0] * self.shape[1]) * 4
        p, h = ctypes.c_void_p(), ctypes.create_string_buffer(64)
        ok = True
        try:
            _cabi.check(_cabi.lib().bff_ipc_alloc(nbytes, ctypes.byref(p), h), "bff_ipc_alloc")
        except _cabi.BffError:
            ok = False
        handles = [None] * ctx.world_size
        dist.all_gather_object(handles, bytes(h.raw) if ok else None)
        self.local_ptr = p.value if ok else None
        self.addrs, self._opened = [], []
        if ok and all(x is not None for x in handles):
            for r, hb in enumerate(handles):
                if r == ctx.rank:
                    self.addrs.append(p.value)
                    continue
                q = ctypes.c_void_p()
                try:
                    _cabi.check(_cabi.lib().bff_ipc_open(ctypes.create_string_buffer(hb, 64), ctypes.byref(q)), "bff_ipc_open")
                except _cabi.BffError:
                    ok = False
                    break
                self._opened.append(q.value)
                self.addrs.append(q.value)
        else:
            ok = False
        flags = [None] * ctx.world_size
        dist.all_gather_object(flags, ok)
        if not all(flags):
            self.close()
            raise _cabi.BffError("CUDA IPC is not available between the ranks: falling back to the NCCL row exchange")
        self.local = torch.as_tensor(_RawCuda(p.value, self.shape, "<i4"), device=device)
        self.flag = torch.zeros((1,), dtype=torch.int32, device=device)

    def close(self):
        import ctypes
        from . import _cabi
        for a in self._opened:
            _cabi.lib().bff_ipc_close(ctypes.c_void_p(a))
        self._opened = []
        if self.local_ptr:
            _cabi.lib().bff_ipc_free(ctypes.c_void_p(self.local_ptr))
            self.local_ptr = None


def gather_final(comm: Comm, plan: ShardPlan, rf, to_host: bool = True) -> dict:
    """Final dict of a point-sharded bff_refine_device call: every rank holds its column slice of the saved rows (the
    decisions, hence the counts, are the same everywhere); the slices are all-gathered and re-assembled into (k, N) bool."""
    from . import ops
    h = rf.header()
    k = h["n_final"]
    if k == 0:
        return {"ins": [], "conf": [], "final_class": []}
    slices = comm.allgather_rows(rf.final_rows[:k])
    rows = ops.rows_unslice(slices, plan.words)
    ins = ops.unpack_bits(rows, plan.n_points)
    return {"ins": ops.to_host(ins) if to_host else ins, "conf": torch.from_numpy(h["final_conf"].copy()), "final_class": [rf.query] * k}


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
