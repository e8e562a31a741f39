"""TEST INFRASTRUCTURE ONLY -- freezes the output of the reference's depth ingest lines
(tools/projection_2d_to_3d.py:431-436: imread(...).astype(float32) / 1000, cv2.resize(depth, (W, H)))
for small seeded uint16 images into tests/golden/resize.npz. cv2.resize is the reference's own call; the
arrays pin oracle.resize_bilinear (and through it the in-kernel sampling) where cv2 itself is not at hand.
Run in the build container: python oracle/make_golden_resize.py  (cv2 4.13.0, IPP build)."""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [((48, 64), (97, 130)), ((97, 130), (48, 64)), ((60, 80), (121, 162)), ((30, 40), (30, 40)), ((33, 47), (64, 95))]


def main():
    out = {}
    for k, (src_hw, dst_hw) in enumerate(CASES):
        rng = np.random.default_rng(500 + k)
        raw = rng.integers(300, 6000, src_hw).astype(np.uint16)
        raw[rng.random(src_hw) < 0.05] = 0                      # sensor holes
        depth = raw.astype(np.float32) / 1000                   # :431-435
        depth = cv2.resize(depth, (dst_hw[1], dst_hw[0]))       # :436
        out[f"raw{k}"], out[f"dst{k}"] = raw, depth
    out["cv2_version"] = np.array(cv2.__version__)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "resize.npz"), **out)
    print({k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
