"""Dev tool: bandwidth of bff_upload (SM-driven zero-copy H2D) vs cudaMemcpyAsync for a 184 MB pinned buffer."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))
import torch
from bff_b200 import ops
dev = torch.device("cuda:0")
src = torch.empty(184 << 20, dtype=torch.uint8).pin_memory()
dst = torch.empty_like(src, device=dev)
def t(fn, n=8):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
ms = t(lambda: dst.copy_(src, non_blocking=True)); print(f"cudaMemcpyAsync: {ms:.3f} ms  {src.numel()/ms/1e6:.1f} GB/s")
ms = t(lambda: ops.upload(dst, src)); print(f"bff_upload     : {ms:.3f} ms  {src.numel()/ms/1e6:.1f} GB/s")
