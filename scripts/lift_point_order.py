"""Dev tool: how much the lift kernel's warp-level culling depends on the spatial coherence of consecutive points.
Times bff_lift_scene on the bench's c2 scene with the cloud in its mesh-like order, sorted along a Morton curve, and
randomly shuffled (the worst case: every warp's bounding sphere spans the room, nothing is culled)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))
import numpy as np, torch
import bench
from bff_b200 import ops

dev = torch.device("cuda:0")
inp = bench.build_inputs(sys.argv[1] if len(sys.argv) > 1 else "c2", 0, 1, dev)
g = bench.GpuScene(inp, dev)
d = g.upload()
masks2d = g.masks2d(d)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
xyz = d["xyz"]

def morton(x):
    q = ((x[:, :3] - x[:, :3].min(0).values) / (x[:, :3].max(0).values - x[:, :3].min(0).values + 1e-9) * 1023).long()
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F; v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)

orders = {"mesh-like (as generated)": torch.arange(g.n, device=dev), "morton sorted": torch.argsort(morton(xyz)),
          "random shuffle": torch.randperm(g.n, device=dev, generator=torch.Generator(device=dev).manual_seed(0))}
ref = None
for name, perm in orders.items():
    d2 = dict(d, xyz=xyz[perm].contiguous())
    ts = []
    for it in range(9):
        flush.fill_(1)
        out = g.lift(d2, masks2d, timing=True)
        torch.cuda.synchronize()
        if it >= 2: ts.append(g.lift_events[0].elapsed_time(g.lift_events[1]))
    chk = (int(out.masked_counts.sum()), int(out.viewed_counts.sum()))
    ref = ref or chk
    print(f"{name:28s} {min(ts):.3f} ms  {'same votes' if chk == ref else 'MISMATCH'}", flush=True)

# what lift.lift does for the shuffled cloud: spread check + Morton sort + launch + un-permute (bff_rows_permute)
from bff_b200 import lift as lift_mod
shuf = xyz[orders["random shuffle"]].contiguous()
def whole():
    s = lift_mod.point_order_spread(shuf)
    perm = lift_mod.morton_order(shuf)
    out = g.lift(dict(d, xyz=shuf[perm].contiguous()), masks2d)
    inv = torch.empty_like(perm); inv[perm] = torch.arange(g.n, device=dev)
    rows = ops.rows_permute(out.rows, inv.to(torch.int32), g.n)
    return rows, out.masked_counts[:, inv], out.viewed_counts[inv]
for _ in range(3): whole()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): whole()
e1.record(); torch.cuda.synchronize()
print(f"shuffled cloud through check + sort + lift + un-permute: {e0.elapsed_time(e1) / 5:.3f} ms (plain launch on it: see 'random shuffle')")
