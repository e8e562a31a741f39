"""CPU: host-side decision logic of the product (no GPU): component search fast path == general BFS,
sequential float32 means, label equality, frame planning, RLE run clipping (Python-slice semantics)."""
import pytest
import numpy as np
import torch

from bff_b200 import aggregate, lift, rle


def test_connected_groups_fast_path_equals_general_bfs():
    rng = np.random.default_rng(0)
    for _ in range(200):
        n = int(rng.integers(1, 70))
        a = rng.random((n, n)) < 0.05
        a = a | a.T
        for i in range(n):
            if rng.random() < 0.3:
                a[i, :] = False
                a[:, i] = False
        assert aggregate.connected_groups(torch.from_numpy(a)) == aggregate._connected_groups_general(a)
    # asymmetric input takes the general path: reachability, not components
    a = np.zeros((4, 4), bool)
    a[0, 1] = a[1, 2] = True
    assert aggregate.connected_groups(torch.from_numpy(a)) == [[1, 2], []]


def test_sequential_float32_mean_matches_python_sum_of_tensors():
    rng = np.random.default_rng(1)
    for _ in range(100):
        n = int(rng.integers(1, 60))
        c = torch.from_numpy(rng.uniform(0.2, 0.45, n).astype(np.float32))
        ref = sum([c[k] for k in range(n)]) / n                       # what merge_masks does (:220-224)
        got = np.cumsum(c.numpy(), dtype=np.float32)[-1] / np.float32(n)
        assert ref.item() == float(got)


def test_label_equality():
    m = aggregate.label_equality(["a", "b", "a", "c c", "b"])
    exp = torch.tensor([[1, 0, 1, 0, 0], [0, 1, 0, 0, 1], [1, 0, 1, 0, 0], [0, 0, 0, 1, 0], [0, 1, 0, 0, 1]], dtype=torch.bool)
    assert torch.equal(m, exp) and aggregate.label_equality([]).shape == (0, 0)


def test_plan_frames_visits_each_frame_once_and_keeps_file_order():
    q = lift.QueryMasks("q", ["30", "10", "50"], [[{}], [{}, {}], [{}]], [None] * 3, [["q"], ["q", "q"], ["q"]])
    order, viewed, per_frame = lift.plan_frames([q], ["0", "10", "20", "30", "40"])
    assert order == ["30", "10", "50", "0", "20", "40"]                # mask frames in FILE order, then view-only frames
    assert viewed == [True, True, False, True, True, True]             # frame 50 has masks but is not a listed view
    assert lift.query_row_index(per_frame, [q]) == [[0, 1, 2, 3]]
    rles, mq, off = lift.mask_layout(per_frame, [q], 0, len(order))
    assert off == [0, 1, 3, 4, 4, 4, 4] and mq == [0, 0, 0, 0]


def test_rle_runs_follow_python_slice_semantics():
    r = {"length": 50, "counts": np.array([0, 5, 48, 10, 60, 3, 10, 0, 20, 5], np.int64)}
    lo, hi, L = rle.runs_of(r)
    m = np.zeros(50, np.uint8)
    for s, n in zip(r["counts"][0::2], r["counts"][1::2]):            # the reference's loop (refinement.py:33-38)
        m[int(s) - 1:int(s) - 1 + int(n)] = 1
    d = np.zeros(50, np.uint8)
    for a, b in zip(lo, hi):
        d[a:b] = 1
    assert np.array_equal(d, m) and L == 50
    runs, length = rle.runs_for_device([r, {"length": 50, "counts": np.zeros(0, np.int64)}])
    assert length == 50 and set(runs[:, 0].tolist()) == {0}


def test_resize_table_reproduces_cv2_resize():
    """bff_resize_table (host function of libbff.so) applied with float32 fmaf blends == oracle.resize_bilinear
    (itself pinned to cv2.resize in test_oracle_golden.py)."""
    import numpy as np
    import oracle as O
    from bff_b200 import ops
    rng = np.random.default_rng(3)
    src = (rng.integers(0, 6000, (48, 64)).astype(np.uint16)).astype(np.float32) / 1000
    for (hd, wd) in [(97, 130), (24, 31), (48, 64)]:
        tx, ty = ops.resize_table(wd, 64), ops.resize_table(hd, 48)
        x0, wx = tx[:, 0], tx[:, 1].copy().view(np.float32)
        y0, wy = ty[:, 0], ty[:, 1].copy().view(np.float32)
        x1, y1 = np.minimum(x0 + 1, 63), np.minimum(y0 + 1, 47)

        def fma(a, b, c):   # exact product and sum in float64, one rounding to float32 (no double rounding at these sizes)
            return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)
        h = fma(src[:, x1] - src[:, x0], np.broadcast_to(wx, (48, wd)), src[:, x0])
        got = fma(h[y1] - h[y0], np.broadcast_to(wy[:, None], (hd, wd)), h[y0])
        np.testing.assert_array_equal(got, O.resize_bilinear(src, hd, wd))


def test_lpt_scene_assignment_and_costs():
    """Scene-parallel scheduling (BASELINE config 5): every scene exactly once, loads balanced like LPT guarantees
    (max load <= mean + largest item), deterministic, sorted per rank; costs read from the golden tree."""
    import os
    import numpy as np
    import yaml
    from conftest import GOLDEN
    from bff_b200 import schedule
    rng = np.random.default_rng(0)
    costs = (rng.uniform(4e4, 4.2e5, 312) * rng.integers(58, 544, 312)).tolist()       # SURVEY F8 ranges
    for world in (1, 2, 8):
        parts = schedule.lpt_assign(costs, world)
        assert sorted(i for p in parts for i in p) == list(range(312)) and all(p == sorted(p) for p in parts)
        loads = [sum(costs[i] for i in p) for p in parts]
        assert max(loads) <= sum(costs) / world + max(costs)
        assert parts == schedule.lpt_assign(costs, world)
        if world == 8:
            assert max(loads) / (sum(costs) / world) < 1.02                              # 312 scenes balance to within 2 %
    assert schedule.lpt_assign([], 4) == [[], [], [], []]
    tree = os.path.join(GOLDEN, "tree")
    cfg = yaml.safe_load(open(os.path.join(tree, "config.yaml")))
    cfg = {k: (os.path.join(tree, v[2:]) if isinstance(v, str) and v.startswith("./") else v) for k, v in cfg.items()}
    assert schedule.scene_cost(cfg, "scene0001_00") == 3000 * 8 and schedule.scene_cost(cfg, "scene0002_00") == 2500 * 6
    assert schedule.dist_mode(3, 8) == "frame" and schedule.dist_mode(312, 8) == "scene"


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the CPU arm the driver runs beside ours): one JSON line with the contract's keys, the
    same metric/unit as the GPU arm, a cpu_baseline describing the run and an e2e object without transfers."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-budget", "5", "--workload", "tiny"], capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "point-views/s" and d["higher_is_better"] is True
    assert d["metric"].startswith("point-view projections/sec") and d["value"] > 0 and d["n_gpus"] == 1
    assert d["steps"] == 1 and d["warmup"] == 0 and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_class_batches_and_orchestrator_checkpoint(tmp_path, monkeypatch):
    """bff_b200/run_classes.py: classes go through both stages in batches, finished stages (process_checkpoint.txt, the
    reference orchestrator's file, evaluation/eval/run_evl.py:28-54) are skipped on a re-run."""
    import yaml
    from bff_b200 import run_classes
    monkeypatch.chdir(tmp_path)
    calls = []
    s2 = lambda a: calls.append(("lift", a[3:])) or 0
    rf = lambda a: calls.append(("refine", a[3:])) or 0
    ck = run_classes.run("c.yaml", ["a", "b", "c"], batch=2, stage2_main=s2, refine_main=rf)
    assert calls == [("lift", ["a", "b"]), ("refine", ["a", "b"]), ("lift", ["c"]), ("refine", ["c"])]
    assert yaml.safe_load(open("process_checkpoint.txt")) == ck == {c: {"projection_2d_to_3d": True, "refinement": True} for c in "abc"}
    calls.clear()
    ck2 = yaml.safe_load(open("process_checkpoint.txt"))
    ck2["b"]["refinement"] = False
    yaml.safe_dump(ck2, open("process_checkpoint.txt", "w"))
    run_classes.run("c.yaml", ["a", "b", "c", "d"], batch=3, stage2_main=s2, refine_main=rf)
    assert calls == [("refine", ["b"]), ("lift", ["d"]), ("refine", ["d"])]


def test_rle_pairs_for_device_keeps_the_wire_format():
    from bff_b200 import rle
    a = {"length": 50, "counts": np.array([3, 4, 10, 2, 7], np.int64)}           # unpaired trailing start
    b = {"length": 50, "counts": np.zeros(0, np.int64)}
    c = {"length": 50, "counts": torch.tensor([1, 50])}
    counts, off, length = rle.pairs_for_device([a, b, c])
    assert length == 50 and counts.dtype == np.int64 and counts.tolist() == [3, 4, 10, 2, 1, 50] and off.tolist() == [0, 2, 2, 3]
    with pytest.raises(ValueError):
        rle.pairs_for_device([a, {"length": 51, "counts": np.zeros(0, np.int64)}])


def test_sweep_plan_covers_every_scene_once_and_balances_the_ranks():
    """bench.py --workload c5 (BASELINE configs[4]): 312 scene shapes in the val set's ranges, bucketed into size classes,
    spread over the ranks by schedule.lpt_assign -- every scene served exactly once, loads within 1 %."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    for world in (1, 2, 8):
        pool, cls, shares = bench.sweep_plan(world)
        assert len(pool) == bench.SWEEP_POOL and len(cls) == bench.SWEEP_SCENES and len(shares) == world
        assert sorted(i for s in shares for i in s) == list(range(bench.SWEEP_SCENES))
        assert all(40_000 <= p <= 420_000 and 58 <= f <= 543 for p, f in pool)
        assert [p * f for p, f in pool] == sorted(p * f for p, f in pool)
        load = [sum(pool[cls[i]][0] * pool[cls[i]][1] for i in s) for s in shares]
        assert max(load) <= 1.01 * min(load)
        assert (pool, cls, shares) == bench.sweep_plan(world)          # every rank computes the same plan


def test_pipeline_row_limits_match_the_header():
    """stage2.py / refine.py route jobs beyond the fused entry points' row limits to the next path down; the numbers they
    compare with are the ones include/bff.h documents and the library enforces."""
    import os
    import re
    from bff_b200 import refine, stage2
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "bff.h")).read().replace(" ", " ")
    num = lambda pat: int(re.search(pat, text).group(1).replace(" ", ""))
    assert stage2.MAX_FUSED_ROWS == num(r"bff_stage2_aggregate: n_rows <= ([\d ]+\d)")
    assert stage2.MAX_DEVICE_ROWS == num(r"bff_stage2_device: n_rows <= ([\d ]+\d)")
    assert refine.MAX_FUSED_ROWS == num(r"bff_refine_match: n2 <= ([\d ]+\d)")
    assert refine.MAX_DEVICE_ROWS == num(r"bff_refine_device: cap2 <= ([\d ]+\d)")
    src = "".join(open(os.path.join(root, "beyond-fixed-forms_b200", "csrc", f)).read() for f in ("devpipe.cu", "pipeline.cu"))
    for limit in (stage2.MAX_FUSED_ROWS, stage2.MAX_DEVICE_ROWS, refine.MAX_FUSED_ROWS, refine.MAX_DEVICE_ROWS):
        assert f"<= {limit}" in src


def test_fast_pose_reader_returns_np_loadtxt_bits(tmp_path):
    """scene_io.read_matrix4 == np.loadtxt on pose files: the golden tree's, random matrices written like ScanNet writes
    them (%.18e and short decimals), invalid poses (-inf), nan, and layouts that must fall back to np.loadtxt."""
    import glob
    import os
    import warnings
    from bff_b200 import scene_io
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    files = sorted(glob.glob(os.path.join(root, "tests", "golden", "tree", "data", "Scannet200", "Scannet200_2D", "*", "pose", "*.txt")))
    assert len(files) >= 10
    rng = np.random.default_rng(3)
    for k in range(40):
        m = rng.normal(size=(4, 4)) * float(rng.choice([1e-6, 1.0, 1e5]))
        path = tmp_path / f"m{k}.txt"
        np.savetxt(path, m, fmt=["%.18e", "%.6f", "%r", "%.3g"][k % 4] if k % 4 != 2 else "%.17g")
        files.append(str(path))
    hostile = {"inf.txt": "-inf -inf -inf -inf\n" * 4, "nan.txt": "nan 1 2 3\n0 1 0 0\n0 0 1 0\n0 0 0 1\n",
               "crlf.txt": "1 0 0 0.5\r\n0 1 0 0\r\n0 0 1 0\r\n0 0 0 1\r\n", "blank.txt": "\n1 0 0 0\n\n0 1 0 0\n0 0 1 0\n0 0 0 1\n\n",
               "comment.txt": "# pose\n1 0 0 0\n0 1 0 0\n0 0 1 0\n0 0 0 1\n", "one_line.txt": "1 0 0 0 0 1 0 0 0 0 1 0 0 0 0 1\n",
               "three.txt": "1 0 0\n0 1 0\n0 0 1\n"}
    for name, text in hostile.items():
        (tmp_path / name).write_text(text)
        files.append(str(tmp_path / name))
    for f in files:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            want = np.loadtxt(f)
            got = scene_io.read_matrix4(f)
        assert got.shape == want.shape and got.dtype == want.dtype, f
        assert np.array_equal(got.view(np.int64), want.view(np.int64)), f


@pytest.mark.parametrize("threads", [1, 8])
def test_load_frames_equals_the_reference_calls_per_frame(monkeypatch, threads):
    """scene_io.load_frames (thread pool for the PNGs, poses parsed beside it) hands over exactly what the reference computes
    per frame: np.linalg.inv(np.loadtxt(pose)) (:422-425) and cv2.imread(depth, -1) (:431)."""
    import glob
    import os
    import cv2
    from bff_b200 import scene_io
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d2 = os.path.join(root, "tests", "golden", "tree", "data", "Scannet200", "Scannet200_2D")
    scene = sorted(os.listdir(d2))[0]
    ids = sorted((os.path.basename(f)[:-4] for f in glob.glob(os.path.join(d2, scene, "pose", "*.txt"))), key=int)
    assert len(ids) >= 5
    first = cv2.imread(os.path.join(d2, scene, "depth", f"{ids[0]}.png"), cv2.IMREAD_UNCHANGED)
    monkeypatch.setattr(scene_io, "INGEST_THREADS", threads)
    fs = scene_io.load_frames(d2, scene, ids, [True] * len(ids), first.shape[1], first.shape[0])
    assert fs.pose_inv.shape == (len(ids), 4, 4) and fs.depth.shape == (len(ids),) + first.shape and fs.depth.dtype == np.uint16
    for i, fid in enumerate(ids):
        want = np.linalg.inv(np.loadtxt(os.path.join(d2, scene, "pose", f"{fid}.txt")))
        assert np.array_equal(fs.pose_inv[i].view(np.int64), want.view(np.int64))
        assert np.array_equal(fs.depth[i], cv2.imread(os.path.join(d2, scene, "depth", f"{fid}.png"), cv2.IMREAD_UNCHANGED))
