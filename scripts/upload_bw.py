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
# both at once: the copy engine moves the head of the buffer while the SMs pull the tail over the same link
sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
for frac in (0.3, 0.4, 0.5, 0.6):
    cut = (int(src.numel() * frac) >> 12) << 12
    def both():
        ev = torch.cuda.Event(); ev.record()
        sa.wait_event(ev); sb.wait_event(ev)
        with torch.cuda.stream(sa):
            dst[:cut].copy_(src[:cut], non_blocking=True)
        ops.upload(dst[cut:], src[cut:], sb)
        torch.cuda.current_stream().wait_stream(sa); torch.cuda.current_stream().wait_stream(sb)
    ms = t(both); print(f"copy engine {frac:.1f} + SM {1-frac:.1f}: {ms:.3f} ms  {src.numel()/ms/1e6:.1f} GB/s")
