"""Dev tool: time bff_mask_iou (symmetric) over a sweep of row counts at c2 width, to spot cliffs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))
import torch
from bff_b200 import ops
dev = torch.device("cuda:0")
words = int(sys.argv[1]) if len(sys.argv) > 1 else 7813
g = torch.Generator(device=dev).manual_seed(1)
for n in (96, 128, 256, 525, 640, 896, 1024, 1031, 1086, 1351, 1445, 2048, 2858, 4299):
    rows = torch.randint(-2**31, 2**31 - 1, (n, words), dtype=torch.int32, device=dev, generator=g)
    rows &= torch.randint(-2**31, 2**31 - 1, (n, words), dtype=torch.int32, device=dev, generator=g)
    for _ in range(2): ops.mask_iou(rows, want_iou=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ops.mask_iou(rows, want_iou=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"n={n:5d} words={words}: {ms:8.3f} ms  {n*n*words*32*2/ms/1e9:8.1f} T int8 op/s (full square)", flush=True)
