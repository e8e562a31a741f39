"""Dev tool: time bff_lift_scene for every experiment build in variants/*.so (see build.py -D... --out=...)
and a sweep of frames_per_chunk, on the bench's c2 scene. Prints ms per launch (L2 flushed, CUDA events
recorded inside the library around the launch)."""
import glob
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "beyond-fixed-forms_b200"))
import torch

import bench
from bff_b200 import _cabi, ops


def main():
    dev = torch.device("cuda:0")
    wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
    inp = bench.build_inputs(wl, 0, 1, dev)
    g = bench.GpuScene(inp, dev)
    d = g.upload()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    libs = [None] + sorted(glob.glob(os.path.join(ROOT, "variants", "*.so")))
    ref = None
    for lib in libs:
        _cabi._lib = None
        _cabi.LIB_PATH = lib or os.path.join(ROOT, "beyond-fixed-forms_b200", "bff_b200", "libbff.so")
        masks2d = g.masks2d(d)
        for fpc in (16, 20, 25, 30, 38, 50):
            ts = []
            for it in range(13):
                flush.fill_(1)
                out = ops.lift_scene(d["xyz"], d["pinv"], inp["K"], d["depth"], g.h, g.w, masks2d=masks2d, mask_offset=d["offsets"],
                                     frame_viewed=d["viewed"], depth_range=(0.0, 65.535), frames_per_chunk=fpc, timing_events=g.lift_events)
                torch.cuda.synchronize()
                if it >= 3:
                    ts.append(g.lift_events[0].elapsed_time(g.lift_events[1]))
            chk = (int(out.masked_counts.sum()), int(out.viewed_counts.sum()), int(out.rows.to(torch.int64).sum()))
            if ref is None:
                ref = chk
            print(f"{os.path.basename(_cabi.LIB_PATH):22s} fpc={fpc:4d}  min {min(ts):.4f}  med {sorted(ts)[len(ts)//2]:.4f} ms  {'ok' if chk == ref else 'MISMATCH'}", flush=True)


if __name__ == "__main__":
    main()
