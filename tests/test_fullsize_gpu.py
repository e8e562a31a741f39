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
    masks2d = g.masks2d(d)
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


# ------------------------------------------------------------------ c3: the c2 scene with 20 concurrent text queries
@pytest.fixture(scope="module")
def c3(cuda_dev):
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops
    inp = bench.build_inputs("c3", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    masks2d = g.masks2d(d)
    return inp, g, d, masks2d


def test_c3_shared_projection_equals_one_launch_per_query(c3):
    """BASELINE configs[2]: 20 queries lifted in ONE launch (shared projection / visibility / viewed counts, per-query
    votes) give, for every query, exactly the rows and votes of a launch that carries only that query's masks --
    which is what one reference process per --cls computes."""
    from bff_b200 import ops
    inp, g, d, masks2d = c3
    assert g.nq == 20 and g.n == 250_000 and g.f == 300 and g.n_masks > 800
    full = g.lift(d, masks2d)
    mq = d["mask_query"].cpu().numpy()
    off = d["offsets"].cpu().numpy()
    pop = ops.rows_and_popcount(full.rows.clone()).cpu().numpy()
    for qi in range(g.nq):                                                   # votes are one per set row bit, per query
        assert int(pop[mq == qi].sum()) == int(full.masked_counts[qi].sum())
    assert int(full.masked_counts.sum()) > 100_000
    busiest = np.argsort(-np.bincount(mq, minlength=g.nq))[:3].tolist()
    for qi in busiest:
        sel = np.nonzero(mq == qi)[0]
        per_frame = np.array([np.count_nonzero((sel >= off[f]) & (sel < off[f + 1])) for f in range(g.f)])
        off_q = torch.from_numpy(np.concatenate([[0], np.cumsum(per_frame)]).astype(np.int32)).to(g.dev)
        sel_t = torch.from_numpy(sel).to(g.dev)
        one = ops.lift_scene(d["xyz"], d["pinv"], inp["K"], d["depth"], g.h, g.w, masks2d=masks2d[sel_t], mask_offset=off_q,
                             frame_viewed=d["viewed"], depth_range=(0.0, 65.535))
        assert len(sel) > 10
        assert torch.equal(one.rows, full.rows[sel_t])
        assert torch.equal(one.masked_counts[0], full.masked_counts[qi])
        assert torch.equal(one.viewed_counts, full.viewed_counts)
    # the launch layout and each query's file order agree with the host bookkeeping
    for qi in range(g.nq):
        assert sorted(g.row_index[qi]) == np.nonzero(mq == qi)[0].tolist()


def test_c3_subset_exact_against_oracle(c3):
    import oracle as O
    inp, g, d, masks2d = c3
    full = g.lift(d, masks2d)
    nf = 60                                                                   # first 60 planned frames, every 97th point
    off = d["offsets"].cpu().numpy()[:nf + 1]
    part = ops_lift_block(c3, 0, nf)
    sel = np.arange(0, g.n, 97)[:2560]
    perm = [inp["frame_ids"].index(fid) for fid in g.frame_ids][:nf]
    depth = (inp["depth"][perm].astype(np.float32) / 1000).astype(np.float32)
    masks = unpack_rows(masks2d[:off[-1]].cpu().numpy(), g.h * g.w).reshape(-1, g.h, g.w).astype(np.uint8)
    rows_o, mc_o, vc_o = O.lift_scene(inp["xyz"][sel], inp["pinv"][perm], inp["K"], depth, 0.08, masks, off,
                                      d["mask_query"].cpu().numpy()[:off[-1]], np.asarray(g.viewed[:nf], np.uint8), n_query=g.nq)
    assert vc_o.sum() > 2_000 and rows_o.sum() > 300
    np.testing.assert_array_equal(unpack_rows(part.rows.cpu().numpy(), g.n)[:, sel], rows_o)
    np.testing.assert_array_equal(part.masked_counts.cpu().numpy()[:, sel], mc_o)
    np.testing.assert_array_equal(part.viewed_counts.cpu().numpy()[sel], vc_o)
    assert torch.equal(part.rows, full.rows[:off[-1]])


def ops_lift_block(fx, lo, hi, fpc=0, depth=None, native="same"):
    """Frames [lo, hi) of a GpuScene fixture (all queries)."""
    from bff_b200 import ops
    inp, g, d, masks2d = fx
    off = d["offsets"][lo:hi + 1]
    m0, m1 = int(off[0]), int(off[-1])
    return ops.lift_scene(d["xyz"], d["pinv"][lo:hi], inp["K"], d["depth"][lo:hi] if depth is None else depth, g.h, g.w,
                          masks2d=masks2d[m0:m1] if m1 > m0 else None, mask_offset=(off - off[0]).contiguous() if m1 > m0 else None,
                          mask_query=d["mask_query"][m0:m1].contiguous() if (m1 > m0 and "mask_query" in d) else None, n_queries=g.nq,
                          frame_viewed=d["viewed"][lo:hi], depth_range=(0.0, 65.535), frames_per_chunk=fpc,
                          native_size=g.native if native == "same" else native)


# ------------------------------------------------------------------ c4: one rank's share of the 2M-point / 2000-frame scene
@pytest.fixture(scope="module")
def c4(cuda_dev):
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import ops
    saved = dict(bench.WORKLOADS["c4"])
    bench.WORKLOADS["c4"]["n_frames"] = 96                                   # a 96-frame block keeps the fixture under a minute
    try:
        inp = bench.build_inputs("c4", 0, 1, cuda_dev)
    finally:
        bench.WORKLOADS["c4"] = saved
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    masks2d = g.masks2d(d)
    return inp, g, d, masks2d


def test_c4_in_kernel_resize_blocks_and_chunks(c4):
    """BASELINE configs[3] shape (2M points, 1296x968 masks over 640x480 uint16 depth): the kernel's own bilinear sampling
    equals lifting on the image bff_depth_resize writes (which equals cv2.resize, test_kernels_gpu.py); frame blocks
    add up (the multi-GPU invariant) and chunking is invisible."""
    from bff_b200 import ops
    inp, g, d, _ = c4
    assert g.n == 2_000_000 and (g.h, g.w) == (968, 1296) and g.native == (480, 640)
    full = ops_lift_block(c4, 0, g.f)
    assert full.rows.shape[0] > 60 and int(full.viewed_counts.sum()) > 5_000_000
    pop = ops.rows_and_popcount(full.rows.clone())
    assert int(pop.sum()) == int(full.masked_counts.sum()) > 50_000
    resized = ops.depth_resize(d["depth"], g.h, g.w)                          # (F, 968, 1296) float32 on the device
    ref = ops_lift_block(c4, 0, g.f, depth=resized, native=None)
    assert torch.equal(ref.rows, full.rows) and torch.equal(ref.masked_counts, full.masked_counts)
    assert torch.equal(ref.viewed_counts, full.viewed_counts)
    del resized, ref
    cut = 41
    a, b = ops_lift_block(c4, 0, cut), ops_lift_block(c4, cut, g.f)
    assert torch.equal(a.masked_counts + b.masked_counts, full.masked_counts)
    assert torch.equal(a.viewed_counts + b.viewed_counts, full.viewed_counts)
    assert torch.equal(torch.cat([a.rows, b.rows]), full.rows)
    o = ops_lift_block(c4, 0, g.f, fpc=96)
    assert torch.equal(o.rows, full.rows) and torch.equal(o.viewed_counts, full.viewed_counts)


def test_c4_subset_exact_against_oracle(c4):
    import oracle as O
    inp, g, d, masks2d = c4
    nf = 24
    part = ops_lift_block(c4, 0, nf)
    off = d["offsets"].cpu().numpy()[:nf + 1]
    sel = np.arange(0, g.n, 389)[:5000]
    perm = [inp["frame_ids"].index(fid) for fid in g.frame_ids][:nf]
    depth = np.stack([O.resize_bilinear(inp["depth"][i].astype(np.float32) / 1000, g.h, g.w) for i in perm])
    masks = unpack_rows(masks2d[:off[-1]].cpu().numpy(), g.h * g.w).reshape(-1, g.h, g.w).astype(np.uint8)
    rows_o, mc_o, vc_o = O.lift_scene(inp["xyz"][sel], inp["pinv"][perm], inp["K"], depth, 0.08, masks, off,
                                      None, np.asarray(g.viewed[:nf], np.uint8))
    assert vc_o.sum() > 2_000 and rows_o.sum() > 50
    np.testing.assert_array_equal(unpack_rows(part.rows.cpu().numpy(), g.n)[:, sel], rows_o)
    np.testing.assert_array_equal(part.masked_counts.cpu().numpy()[0][sel], mc_o[0])
    np.testing.assert_array_equal(part.viewed_counts.cpu().numpy()[sel], vc_o)


def _same_final(a, b, what):
    assert list(a["final_class"]) == list(b["final_class"]), what
    if isinstance(a["ins"], list) or isinstance(b["ins"], list):
        assert isinstance(a["ins"], list) and isinstance(b["ins"], list), what
        return
    assert a["ins"].shape == b["ins"].shape, what
    assert torch.equal(a["ins"].cpu().bool(), b["ins"].cpu().bool()), what
    np.testing.assert_allclose(torch.as_tensor(a["conf"]).float().numpy(), torch.as_tensor(b["conf"]).float().numpy(), rtol=1e-5, err_msg=what)


def test_c2_whole_step_equals_the_cpu_port(c2):
    """The WHOLE timed step at its real size -- lift -> aggregate (525 lifted masks: the tcgen05 intersections, the 525-node
    component search) -> point filter -> overlap resolution -> refinement against 200 stage-1 masks -- with the decisions
    taken on the device (what bench.py times), with the decisions taken on the host, and by the dense CPU port of the
    reference (oracle/reference_port.py, ~1/2 minute): identical final masks, classes, and confidences within 1e-5."""
    sys.path.insert(0, ROOT)
    import bench
    inp, g, d, _ = c2
    g.device_decisions = True
    dev_final = g.step(d)
    dev_s2 = g.last_stage2_dev.saved_dict()[0]
    g.device_decisions = False
    host_final = g.step(d)
    host_s2 = g.last_stage2
    g.device_decisions = True
    assert dev_s2["ins"].shape[0] >= 5 and dev_final["ins"].shape[0] >= 5
    _same_final(dev_s2, host_s2, "stage 2: device vs host decisions")
    _same_final(dev_final, host_final, "final: device vs host decisions")
    torch.set_num_threads(1)                                  # the port's tiny dgemms are fastest single-threaded
    mask_frames, view_frames = bench.cpu_frames_of(inp, inp["depth"].shape[0])
    port = bench.cpu_step(inp, mask_frames, view_frames)[0]
    _same_final(dev_final, port, "final: device decisions vs the CPU port of the reference")


def test_c3_subset_whole_step_equals_the_cpu_port(cuda_dev):
    """Four of c3's twenty concurrent queries over the c2-sized scene (one shared lifting launch, per-query votes and rows):
    device decisions vs the CPU port, which runs the whole path once per query like one reference process per --cls."""
    sys.path.insert(0, ROOT)
    import bench
    old = dict(bench.WORKLOADS["c3"])
    bench.WORKLOADS["c3q4"] = dict(old, n_queries=4, n_frames=120)
    try:
        inp = bench.build_inputs("c3q4", 0, 1, cuda_dev)
    finally:
        del bench.WORKLOADS["c3q4"]
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    g.step(d)
    finals = list(g.last_finals)
    assert len(finals) == 4 and sum(isinstance(f["ins"], torch.Tensor) and f["ins"].shape[0] >= 2 for f in finals) >= 3
    torch.set_num_threads(1)
    mask_frames, view_frames = bench.cpu_frames_of(inp, inp["depth"].shape[0])
    port = bench.cpu_step(inp, mask_frames, view_frames)
    for qi in range(4):
        _same_final(finals[qi], port[qi], f"query {qi}")
