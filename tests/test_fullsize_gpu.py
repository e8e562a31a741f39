"""GPU, BASELINE.json full size (c2: 250k points x 300 frames 640x480): the oracle cannot run the whole
scene in seconds, so parity is checked (a) exactly against the C oracle on a strided subset of the points
over ALL frames, and (b) through size-independent properties of the whole launch: votes are the column sums
of the lifted rows, frame blocks add up (the invariant the multi-GPU all-reduce relies on), chunking does
not change anything, and the IoU matrix is symmetric with popcounts on its diagonal."""
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import unpack_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2(cuda_dev):
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops
    inp = bench.build_inputs("c2", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    masks2d = ops.rle_to_bits_dev(d["runs"], g.n_masks, g.h * g.w)
    return inp, g, d, masks2d


def _lift(c2, lo=None, hi=None, fpc=0, xyz=None):
    from bff_b200 import ops
    inp, g, d, masks2d = c2
    lo, hi = (0, g.f) if lo is None else (lo, hi)
    off = d["offsets"][lo:hi + 1]
    m0, m1 = int(off[0]), int(off[-1])
    return ops.lift_scene(d["xyz"] if xyz is None else xyz, d["pinv"][lo:hi], inp["K"], d["depth"][lo:hi], g.h, g.w,
                          masks2d=masks2d[m0:m1] if m1 > m0 else None, mask_offset=(off - off[0]).contiguous() if m1 > m0 else None,
                          frame_viewed=d["viewed"][lo:hi], depth_range=(0.0, 65.535), frames_per_chunk=fpc)


def test_c2_properties(c2):
    inp, g, d, _ = c2
    from bff_b200 import ops
    full = _lift(c2)
    assert g.n == 250_000 and g.f == 300 and full.rows.shape[0] > 300
    pop = ops.rows_and_popcount(full.rows.clone())
    assert int(pop.sum()) == int(full.masked_counts.sum()) > 100_000          # votes are one per set row bit (:459-461)
    assert int(full.viewed_counts.max()) <= g.f and int(full.viewed_counts.sum()) > 1_000_000
    # frame blocks add up and their rows concatenate (what the frame-sharded path exchanges)
    cut = 137
    a, b = _lift(c2, 0, cut), _lift(c2, cut, g.f)
    assert torch.equal(a.masked_counts + b.masked_counts, full.masked_counts)
    assert torch.equal(a.viewed_counts + b.viewed_counts, full.viewed_counts)
    assert torch.equal(torch.cat([a.rows, b.rows]), full.rows)
    # chunking of the launch is invisible
    for fpc in (7, 64, 256):
        o = _lift(c2, fpc=fpc)
        assert torch.equal(o.rows, full.rows) and torch.equal(o.masked_counts, full.masked_counts)
        assert torch.equal(o.viewed_counts, full.viewed_counts)
    # IoU: symmetric, diagonal intersections are the row popcounts
    inter, pa, _, iou = ops.mask_iou(full.rows)
    assert torch.equal(inter, inter.T) and torch.equal(inter.diagonal(), pa) and torch.equal(pa, pop)
    assert torch.equal(torch.nan_to_num(iou, nan=-1.0), torch.nan_to_num(iou.T, nan=-1.0))


def test_c2_subset_exact_against_oracle(c2):
    import oracle as O
    inp, g, d, masks2d = c2
    full = _lift(c2)
    sel = np.arange(0, g.n, 97)[:2560]                                        # strided subset, all 300 frames
    xyz = inp["xyz"][sel]
    perm = [inp["frame_ids"].index(fid) for fid in g.frame_ids]
    depth = (inp["depth"][perm].astype(np.float32) / 1000).astype(np.float32)
    masks = unpack_rows(masks2d.cpu().numpy(), g.h * g.w).reshape(-1, g.h, g.w).astype(np.uint8)
    rows_o, mc_o, vc_o = O.lift_scene(xyz, inp["pinv"][perm], inp["K"], depth, 0.08, masks, d["offsets"].cpu().numpy(),
                                      None, np.asarray(g.viewed, np.uint8))
    got_rows = unpack_rows(full.rows.cpu().numpy(), g.n)[:, sel]
    assert vc_o.sum() > 10_000 and rows_o.sum() > 1_000
    np.testing.assert_array_equal(got_rows, rows_o)
    np.testing.assert_array_equal(full.masked_counts.cpu().numpy()[0][sel], mc_o[0])
    np.testing.assert_array_equal(full.viewed_counts.cpu().numpy()[sel], vc_o)
