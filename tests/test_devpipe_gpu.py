"""GPU: the device-resident pipelines (bff_stage2_device, bff_refine_device: every decision a kernel, no host round
trip) against the host-decision pipelines (bff_stage2_aggregate / bff_refine_match and the step-by-step torch path, which
the golden trees pin to the reference) and against the dense CPU oracle port -- on randomized mask sets that provoke
merges, chains, ties, empty rows, duplicate rows, NaN IoUs and every filter mode, and on whole synthetic scenes."""
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import pack_rows

pytestmark = pytest.mark.gpu

CFG = {"iou_thres": 0.2, "min_aggragated_masks": 2, "if_occurance_threshold": False, "occurance_threshold": 0.3,
       "if_detected_ratio_threshold": True, "detected_ratio_threshold": 0.38, "remove_filtered_masks": 0.4, "remove_small_masks": 5,
       "stage1_iou_thres": 0.1, "refinment_sim_percentile": 0.2, "refiment_iou_thres": 0.45}


def _object_rows(rng, n_rows, n, n_obj, noise=0.08, empty=2, dup=2, span=(40, 6)):
    """Rows that are noisy copies of a few 'objects' (unions of intervals), so that IoUs spread over (0, 1]; plus empty
    rows (NaN diagonal) and exact duplicates (IoU 1, ties)."""
    objs = np.zeros((n_obj, n), bool)
    for o in objs:
        for _ in range(rng.integers(1, 4)):
            a = rng.integers(0, n)
            o[a:a + rng.integers(n // span[0], n // span[1])] = True
    rows = objs[rng.integers(0, n_obj, n_rows)].copy()
    rows ^= rng.random(rows.shape) < noise * rng.random((n_rows, 1))
    rows &= rng.random(rows.shape) < rng.uniform(0.3, 1.0, (n_rows, 1))          # partial views of an object
    for _ in range(empty):
        rows[rng.integers(0, n_rows)] = False
    for _ in range(dup):
        rows[rng.integers(0, n_rows)] = rows[rng.integers(0, n_rows)]
    return rows


def _same(a, b, what, exact_conf=True):
    assert list(a["final_class"]) == list(b["final_class"]), what
    if isinstance(a["ins"], list) or isinstance(b["ins"], list):
        assert isinstance(a["ins"], list) and isinstance(b["ins"], list) and len(a["ins"]) == len(b["ins"]) == 0, what
        return
    assert a["ins"].shape == b["ins"].shape, what
    if a["ins"].numel():
        assert torch.equal(a["ins"].cpu().bool(), b["ins"].cpu().bool()), what
    ca, cb = torch.as_tensor(a["conf"]).float(), torch.as_tensor(b["conf"]).float()
    if exact_conf:
        assert torch.equal(ca, cb), what
    else:       # torch's mean() of a few float32 values is not the sequential sum (ATen's 4-accumulator / 8-lane order)
        np.testing.assert_allclose(ca.numpy(), cb.numpy(), rtol=1e-6, err_msg=what)


@pytest.mark.parametrize("seed", list(range(12)))
def test_stage2_device_equals_host_decisions_on_random_masks(cuda_dev, seed):
    from bff_b200 import devpipe, stage2
    rng = np.random.default_rng(100 + seed)
    n = int(rng.choice([777, 4096, 20011]))
    n_rows = int(rng.choice([1, 3, 40, 130, 300]))
    rows = _object_rows(rng, n_rows, n, int(rng.integers(2, 12)), empty=min(2, n_rows - 1), dup=min(2, n_rows - 1))
    labels = [str(x) for x in rng.choice(["chair", "office chair", "table"], n_rows, p=[0.7, 0.2, 0.1])]
    if seed % 3 == 0:
        labels = ["chair"] * n_rows
    conf = torch.from_numpy(rng.uniform(0.2, 0.45, n_rows).astype(np.float32))
    mc = torch.from_numpy(rows.sum(0).astype(np.int32))
    vc = torch.from_numpy((rows.sum(0) + rng.integers(0, 9, n)).astype(np.int32))
    cfg = dict(CFG)
    cfg["min_aggragated_masks"] = int(rng.choice([1, 2, 3]))
    mode = seed % 4
    if mode == 1:
        cfg.update(if_occurance_threshold=True, if_detected_ratio_threshold=False)
    elif mode == 2:
        cfg.update(if_occurance_threshold=False, if_detected_ratio_threshold=False)
    if seed % 5 == 0:
        cfg.update(remove_small_masks=5.5, remove_filtered_masks=1)       # float / int scalars: the other promotion branches
    d_rows = torch.from_numpy(pack_rows(rows).view(np.int32)).to(cuda_dev)
    d_mc, d_vc = mc.to(cuda_dev), vc.to(cuda_dev)
    want, reached = stage2.aggregate_and_filter(d_rows, conf, labels, d_mc, d_vc, n, cfg)
    stage2.FUSED = False
    try:
        want_py, _ = stage2.aggregate_and_filter(d_rows, conf, labels, d_mc, d_vc, n, cfg)
    finally:
        stage2.FUSED = True
    s2 = devpipe.stage2_device(d_rows, conf.to(cuda_dev), labels, d_mc, d_vc, n, cfg, mc_cap=int(mc.max()), vc_cap=int(vc.max()))
    got, got_reached = s2.saved_dict()
    _same(want, want_py, "host fused vs step-by-step")
    _same(got, want, f"device vs host decisions (seed {seed}, {n_rows} rows, mode {mode})")
    assert got_reached == reached
    h = s2.header()
    assert all(m == sorted(m) and len(m) >= cfg["min_aggragated_masks"] for m in h["groups"])
    assert [g[0] for g in h["groups"]] == sorted(g[0] for g in h["groups"])


@pytest.mark.parametrize("n_rows", [600, 1100])
def test_stage2_device_many_rows_takes_the_tensor_core_and_global_histogram_paths(cuda_dev, n_rows):
    """>= 96 rows: the all-pairs intersections inside the pipeline run on tcgen05; vote counts up to a few hundred in both
    axes push the (mc, vc) histogram past the shared-memory bins into the global-atomic branch."""
    from bff_b200 import devpipe, stage2
    rng = np.random.default_rng(n_rows)
    n = 50021
    rows = _object_rows(rng, n_rows, n, 40, noise=0.02, empty=3, dup=3, span=(400, 150))
    labels = ["chair"] * n_rows
    conf = torch.from_numpy(rng.uniform(0.2, 0.45, n_rows).astype(np.float32))
    mc = torch.from_numpy(rows.sum(0).astype(np.int32))
    vc = rows.sum(0) + rng.integers(0, 9, n)
    vc[rng.integers(0, n, 200)] += rng.integers(300, 900, 200)                # a few points seen very often: a wide histogram
    vc = torch.from_numpy(vc.astype(np.int32))
    assert (int(mc.max()) + 1) * (int(vc.max()) + 1) > 8192
    d_rows = torch.from_numpy(pack_rows(rows).view(np.int32)).to(cuda_dev)
    want, _ = stage2.aggregate_and_filter(d_rows, conf, labels, mc.to(cuda_dev), vc.to(cuda_dev), n, CFG)
    s2 = devpipe.stage2_device(d_rows, conf.to(cuda_dev), labels, mc.to(cuda_dev), vc.to(cuda_dev), n, CFG, mc_cap=n_rows, vc_cap=int(vc.max()))
    got, _ = s2.saved_dict()
    assert got["ins"].shape[0] >= 2
    _same(got, want, "device vs host decisions")


def _refine_case(rng, n, n1, n2):
    objs = _object_rows(rng, 8, n, 8, noise=0.0, empty=0, dup=0)
    s1 = objs[rng.integers(0, 8, n1)].copy()
    s1 &= rng.random(s1.shape) < rng.uniform(0.4, 1.0, (n1, 1))
    s1[rng.integers(0, n1)] = False
    s2 = objs[rng.integers(0, 8, max(n2, 1))].copy()[:n2]
    s2 &= rng.random(s2.shape) < rng.uniform(0.2, 1.0, (n2, 1))
    if n2 > 2:
        s2[rng.integers(0, n2)] = False                                       # NaN column: argmax picks the first NaN
        s2[rng.integers(0, n2)] = s2[rng.integers(0, n2)]
    return s1, s2


@pytest.mark.parametrize("seed", list(range(14)))
def test_refine_device_equals_host_decisions_on_random_masks(cuda_dev, seed):
    sys.path.insert(0, ROOT)
    import bench
    from bff_b200 import devpipe, refine
    rng = np.random.default_rng(500 + seed)
    n = int(rng.choice([999, 8192, 30011]))
    n1 = int(rng.choice([1, 7, 60, 200]))
    n2 = int(rng.choice([0, 1, 2, 9, 33, 70]))
    s1, s2 = _refine_case(rng, n, n1, n2)
    names = refine.STAGE1_CLASS_NAMES
    query = "office chair"
    cls_idx = rng.choice([names.index("office_chair"), names.index("chair"), names.index("table"), 5, 9], n1)
    stage1 = {"ins": [{"length": n}] * n1, "final_class": torch.from_numpy(cls_idx.astype(np.int64))}
    labels = [names[i] for i in cls_idx]
    conf2 = torch.from_numpy(rng.uniform(0.2, 0.45, n2).astype(np.float32))
    cfg = dict(CFG)
    cfg["stage1_iou_thres"] = float(rng.choice([0.1, 0.02, 0.5]))
    cfg["refiment_iou_thres"] = float(rng.choice([0.45, 0.1]))
    cfg["refinment_sim_percentile"] = float(rng.choice([0.2, 0.0, 0.7]))
    d_s1 = torch.from_numpy(pack_rows(s1).view(np.int32)).to(cuda_dev)
    d_s2 = torch.from_numpy(pack_rows(s2).view(np.int32)).to(cuda_dev) if n2 else torch.zeros((0, (n + 31) // 32), dtype=torch.int32, device=cuda_dev)
    stage2_dict = {"rows": d_s2, "conf": conf2, "ins": torch.zeros((n2, n), dtype=torch.bool)}
    # host decisions (bff_refine_match) and the step-by-step torch path
    res = {}
    for fused in (True, False):
        refine.FUSED = fused
        try:
            p1 = refine.pass1_scene(stage1, stage2_dict, query, cfg["stage1_iou_thres"], bench.fake_similarity, cuda_dev, s1_rows=d_s1.clone())
        finally:
            refine.FUSED = True
        thr = refine.similarity_threshold([p1["sims"]], cfg["refinment_sim_percentile"]) if len(p1["sims"]) else 0.0
        res[fused] = (p1, thr, refine.pass2_scene(p1, query, cfg["refiment_iou_thres"], thr))
    _same(res[True][2], res[False][2], "host fused vs step-by-step", exact_conf=False)
    # device decisions: stage-2 rows in a buffer with spare capacity, the count on the device
    cap2 = n2 + int(rng.integers(1, 5))
    buf = torch.zeros((cap2, d_s1.shape[1]), dtype=torch.int32, device=cuda_dev)
    buf[:n2] = d_s2
    cbuf = torch.zeros((cap2,), dtype=torch.float32, device=cuda_dev)
    cbuf[:n2] = conf2.to(cuda_dev)
    isq, sim = devpipe.stage1_tables(labels, query, bench.fake_similarity, cuda_dev)
    n2_dev = torch.tensor([n2], dtype=torch.int32, device=cuda_dev)
    rf = devpipe.refine_device(d_s1.clone(), buf, n2_dev if seed % 2 else n2, cbuf, isq, sim, n, cfg, query, select_mode=1)
    rf.enqueue_unpack(cap_rows=3 if seed % 3 == 0 else 64)                # small: exercises the overflow fallback
    got = rf.final()
    p1, thr, want = res[True]
    h = rf.header()
    assert h["n2_in"] == n2 and h["n2_out"] == len(p1["sims"])
    assert h["sims"].tolist() == [float(x) for x in p1["sims"]]
    if len(p1["sims"]):
        assert h["sim_thres"] == thr
        assert np.array_equal(h["match_iou"].view(np.int32), torch.as_tensor(p1["ious"]).numpy().view(np.int32))
    _same(got, want, f"device vs host decisions (seed {seed}, n1 {n1}, n2 {n2})")
    # pass 1 only, then pass 2 with a threshold the caller picked (the class-global one of refinement.py:317-324)
    rf0 = devpipe.refine_device(d_s1.clone(), buf, n2, cbuf, isq, sim, n, cfg, query, select_mode=0)
    other_thr = 0.73
    rf0.select(other_thr)
    _same(rf0.final(), refine.pass2_scene(p1, query, cfg["refiment_iou_thres"], other_thr), "select with a caller's threshold")


@pytest.mark.parametrize("occurrence,seed", [(False, 2000), (True, 2000), (False, 2011), (True, 2012), (False, 2013), (False, 2014)])
def test_device_step_equals_host_step_and_port(cuda_dev, occurrence, seed):
    """Whole scenes (lift -> aggregate -> filter -> refine): the device-resident chain, the host-decision chain and the
    dense CPU port of the reference give the same final masks."""
    sys.path.insert(0, ROOT)
    import bench
    inp = bench.build_inputs("tiny", 0, 1, cuda_dev, seed=seed)
    cfg = dict(inp["cfg"])
    if occurrence:
        cfg.update(if_occurance_threshold=True, if_detected_ratio_threshold=False)
    inp["cfg"] = cfg
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    g.device_decisions = True
    dev_final = g.step(d)
    dev_s2 = g.last_stage2_dev.saved_dict()[0]
    g.device_decisions = False
    host_final = g.step(d)
    _same(dev_s2, g.last_stage2, "stage 2")
    _same(dev_final, host_final, "final")
    assert dev_final["ins"].shape[0] >= 1
    mask_frames, view_frames = bench.cpu_frames_of(inp, inp["depth"].shape[0])
    port = bench.cpu_step(inp, mask_frames, view_frames)[0]
    assert torch.equal(dev_final["ins"].cpu(), port["ins"]) and list(dev_final["final_class"]) == list(port["final_class"])
    np.testing.assert_allclose(dev_final["conf"].numpy(), port["conf"].numpy(), rtol=1e-5)


def test_device_step_multi_query_equals_host_step(cuda_dev):
    sys.path.insert(0, ROOT)
    import bench
    inp = bench.build_inputs("tiny3", 0, 1, cuda_dev)
    g = bench.GpuScene(inp, cuda_dev)
    d = g.upload()
    g.device_decisions = True
    g.step(d)
    dev = list(g.last_finals)
    g.device_decisions = False
    g.step(d)
    host = list(g.last_finals)
    assert len(dev) == len(host) == 3
    for qi in range(3):
        _same(dev[qi], host[qi], f"query {qi}")


def test_unpack_dc_guard_bands(cuda_dev):
    """bff_unpack_bits_dc writes rows [0, min(count, cap)) and nothing else: poisoned bytes before, between and after."""
    from bff_b200 import _cabi
    rng = np.random.default_rng(9)
    for n, cap, count in [(77, 5, 3), (4096, 4, 9), (1000, 6, 0), (33, 3, 3), (250007, 3, 2)]:
        rows = rng.random((max(cap, count), n)) < 0.4
        packed = torch.from_numpy(pack_rows(rows).view(np.int32)).to(cuda_dev)
        guard = 64
        buf = torch.full((guard + cap * n + guard,), 0xAB, dtype=torch.uint8, device=cuda_dev)
        cnt = torch.tensor([count], dtype=torch.int32, device=cuda_dev)
        out = buf[guard:guard + cap * n]
        _cabi.check(_cabi.lib().bff_unpack_bits_dc(_cabi.ptr(packed), _cabi.ptr(cnt), cap, n, _cabi.ptr(out), _cabi.current_stream()), "unpack_dc")
        k = min(cap, count)
        got = buf.cpu().numpy()
        assert (got[:guard] == 0xAB).all() and (got[guard + cap * n:] == 0xAB).all()
        assert np.array_equal(got[guard:guard + k * n].reshape(k, n).astype(bool), rows[:k])
        assert (got[guard + k * n:guard + cap * n] == 0xAB).all()


def test_workspace_arenas_survive_growing_jobs(cuda_dev):
    """A sweep of scenes with growing mask counts (the n^2 intersection scratch grows with every new record size) through
    both pipelines: the arenas merge their blocks between calls instead of running into their block limit."""
    from bff_b200 import devpipe, stage2
    rng = np.random.default_rng(5)
    n = 4099
    for k, n_rows in enumerate([5, 40, 90, 150, 260, 410, 640, 900, 1300, 1700] * 4):
        rows = _object_rows(rng, n_rows, n, 12, empty=1, dup=1)
        labels = ["chair"] * n_rows
        conf = torch.from_numpy(rng.uniform(0.2, 0.45, n_rows).astype(np.float32))
        mc = torch.from_numpy(rows.sum(0).astype(np.int32)).to(cuda_dev)
        vc = (mc + 3).contiguous()
        d_rows = torch.from_numpy(pack_rows(rows).view(np.int32)).to(cuda_dev)
        stage2.DEVICE = False
        try:
            want, _ = stage2.aggregate_and_filter(d_rows, conf, labels, mc, vc, n, CFG)
        finally:
            stage2.DEVICE = True
        if k % 7 == 0:
            got, _ = devpipe.stage2_device(d_rows, conf.to(cuda_dev), labels, mc, vc, n, CFG, mc_cap=n_rows, vc_cap=n_rows + 3).saved_dict()
            _same(got, want, f"{n_rows} rows")
