"""TEST INFRASTRUCTURE ONLY -- times the UNMODIFIED reference CLIs (tools/projection_2d_to_3d.py, tools/refinement.py of
/root/reference, run through oracle/run_reference.py) on an on-disk tree of BASELINE configs[0] shape (c1: 150k points,
100 frames 640x480, 1 query, 50 stage-1 masks), next to the oracle port on the same files, and checks that both write
the same masks. The reference tree exists only in the build container, so this cannot run on the GPU box: its output
(profiles/r02_reference_cli_c1_container.json) is the committed record of what the reference itself costs per scene.

    python oracle/time_reference_cli.py [n_scenes] [out.json]
"""
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "beyond-fixed-forms_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)

import torch
import yaml


def main():
    import bench
    import run_reference as RR
    from bff_b200 import synth
    n_scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    out_path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r02_reference_cli_c1_container.json")
    w = bench.WORKLOADS["c1"]
    root = tempfile.mkdtemp(prefix="bff_refcli_")
    scenes = {}
    for i in range(n_scenes):
        spec = synth.SceneSpec(n_points=w["n_points"], n_frames=w["n_frames"], width=w["width"], height=w["height"], n_boxes=w["n_boxes"],
                               n_stage1=w["n_stage1"], seed=3000 + i, n_queries=1)
        scenes[f"scene{i + 1:04d}_00"] = synth.make_scene(spec)
    cfg_path = synth.write_tree(root, scenes, [bench.QUERY], device="cpu")
    cfg = yaml.safe_load(open(cfg_path))
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    res = {"workload": f"c1: {n_scenes} scene(s) x {w['n_points']} points x {w['n_frames']} frames 640x480, 1 query, on-disk tree",
           "host": f"build container, {threads} cores, torch {torch.__version__}", "reference_commit": "1d3aff0f"}
    try:
        t0 = time.perf_counter()
        RR.run_cli("projection_2d_to_3d.py", cfg_path, bench.QUERY, root)
        t1 = time.perf_counter()
        RR.run_cli("refinement.py", cfg_path, bench.QUERY, root)
        t2 = time.perf_counter()
        ref = {}
        for sid in scenes:
            ref[sid] = (torch.load(os.path.join(cfg["mask_3d_dir"], bench.QUERY, f"{sid}.pth"), weights_only=False),
                        torch.load(os.path.join(cfg["final_output_dir"], bench.QUERY, f"{sid}.pth"), weights_only=False))
        shutil.rmtree(cfg["mask_3d_dir"]); shutil.rmtree(cfg["final_output_dir"])
        os.makedirs(cfg["mask_3d_dir"]); os.makedirs(cfg["final_output_dir"])
        # the oracle port on the same files
        import helpers
        with RR.reference_env(root):
            pass                                    # (installs the fake clip module both sides share)
        helpers.fake_clip_sim = lambda q, l: float((lambda a, b: (a @ b.T / (a.norm(dim=-1, keepdim=True) * b.norm(dim=-1, keepdim=True).T)).item())(
            RR.fake_clip_similarity_vector(q), RR.fake_clip_similarity_vector(l)))
        cwd = os.getcwd()
        os.chdir(root)
        t3 = time.perf_counter()
        stage2, finals = helpers.port_run_class(cfg, bench.QUERY)
        t4 = time.perf_counter()
        os.chdir(cwd)
        same = True
        for sid in scenes:
            r2, rf = ref[sid]
            same &= isinstance(rf["ins"], torch.Tensor) == isinstance(finals[sid]["ins"], torch.Tensor)
            if isinstance(rf["ins"], torch.Tensor):
                same &= bool(torch.equal(rf["ins"].cpu(), finals[sid]["ins"])) and list(rf["final_class"]) == list(finals[sid]["final_class"])
            same &= bool(torch.equal(torch.as_tensor(r2["ins"]).cpu().bool(), torch.as_tensor(stage2[sid]["ins"]).bool())) if isinstance(r2["ins"], torch.Tensor) and r2["ins"].numel() else True
        pv = n_scenes * w["n_points"] * w["n_frames"]
        res.update({"reference_seconds": {"projection_2d_to_3d.py": t1 - t0, "refinement.py": t2 - t1, "per_scene": (t2 - t0) / n_scenes},
                    "reference_point_views_per_s": pv / (t2 - t0),
                    "port_seconds": {"both_stages": t4 - t3, "per_scene": (t4 - t3) / n_scenes}, "port_point_views_per_s": pv / (t4 - t3),
                    "lifted_masks": {sid: int(torch.as_tensor(ref[sid][0]["ins"]).shape[0]) for sid in scenes},
                    "port_masks_identical_to_reference": bool(same)})
    finally:
        shutil.rmtree(root, ignore_errors=True)
    json.dump(res, open(out_path, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
