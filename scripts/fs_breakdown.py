"""Dev tool (torchrun, N ranks): where the frame-sharded step's aggregation time goes -- exchange, sharded all-pairs IoU,
the fused aggregate call, refinement. Usage: torchrun --nproc-per-node N scripts/fs_breakdown.py [workload]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))
import numpy as np, torch, torch.distributed as dist
import bench
from bff_b200 import dist as dist_mod, ops, stage2

ctx = dist_mod.init_from_env("nccl")
dev = torch.device("cuda", ctx.local_rank); torch.cuda.set_device(dev)
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
inp = bench.build_inputs(wl, ctx.rank, ctx.world_size, dev)
g = bench.GpuScene(inp, dev, ctx)
gathered = [None] * ctx.world_size
dist.all_gather_object(gathered, (g.n_masks, [float(c) for c in g.conf_q[0]], g.labels_q[0]))
rpr = [x[0] for x in gathered]
conf = torch.tensor([c for x in gathered for c in x[1]], dtype=torch.float32)
labels = [l for x in gathered for l in x[2]]
d = g.upload()
out = g.lift(d, g.masks2d(d))
def T(fn, n=3):
    fn(); torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    for _ in range(n): r = fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3, r
t_ex, (rows, mc, vc) = T(lambda: dist_mod.exchange(ctx, out.rows, rpr, out.masked_counts, out.viewed_counts))
t_pairs, pairs = T(lambda: dist_mod.sharded_pairs(ctx, rows))
t_iou_block, _ = T(lambda: ops.mask_iou(rows[: rows.shape[0] // ctx.world_size], rows, want_iou=False))
t_agg, _ = T(lambda: stage2.aggregate_and_filter(rows, conf, labels, mc[0], vc, g.n, inp["cfg"], keep_rows=True, to_host=False, pairs=pairs))
t_agg0, _ = T(lambda: stage2.aggregate_and_filter(rows, conf, labels, mc[0], vc, g.n, inp["cfg"], keep_rows=True, to_host=False))
if ctx.rank == 0:
    print(f"{wl} world {ctx.world_size} rows {rows.shape[0]}: exchange {t_ex:.2f} ms, sharded_pairs {t_pairs:.2f} ms (one 1/W row block alone {t_iou_block:.2f}), "
          f"aggregate with pairs {t_agg:.2f} ms, aggregate computing pairs itself {t_agg0:.2f} ms", flush=True)
ctx.finalize()
