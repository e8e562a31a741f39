"""Dev tool: bandwidth of bff_upload (SM-driven zero-copy H2D) vs cudaMemcpyAsync for a 184 MB pinned buffer,
over the kernel's grid/unroll knobs (BFF_UPLOAD_BLOCKS / BFF_UPLOAD_UNROLL are read once per process: run per setting)."""
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
tag = f"lib={os.path.basename(os.environ.get('BFF_LIB_PATH', 'libbff.so'))} blocks={os.environ.get('BFF_UPLOAD_BLOCKS', '-')} unroll={os.environ.get('BFF_UPLOAD_UNROLL', '-')}"
ms = t(lambda: dst.copy_(src, non_blocking=True)); ce = src.numel() / ms / 1e6
ms = t(lambda: ops.upload(dst, src)); print(f"{tag}: cudaMemcpyAsync {ce:.1f} GB/s, bff_upload {src.numel()/ms/1e6:.1f} GB/s", flush=True)
