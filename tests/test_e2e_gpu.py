"""GPU: the drop-in CLIs (tools/projection_2d_to_3d.py, tools/refinement.py -> bff_b200.stage2/refine.main)
run on the committed input tree must write the same .pth files as the UNMODIFIED reference CLIs did
(tests/golden/e2e.pth, made by oracle/make_golden.py): instance masks bit-exact, final_class equal,
confidences within 1e-5 relative, empty forms identical."""
import os

import pytest
import torch

from helpers import check_against_expected, copy_tree, load_expected

pytestmark = pytest.mark.gpu


def _install_fake_clip():
    # CLIP is an input of the path; the same deterministic fake the goldens were made with
    import run_reference as RR
    RR._install_stubs()


@pytest.mark.parametrize("cfg_name,out2,outf", [("config.yaml", "output/mask_3d", "output/final_output"),
                                                 ("config_occ.yaml", "output_occ/mask_3d", "output_occ/final_output")])
@pytest.mark.parametrize("cls", ["office chair", "table"])
def test_cli_outputs_match_reference(cuda_dev, tmp_path, monkeypatch, cfg_name, out2, outf, cls):
    from bff_b200 import refine, stage2
    tree = copy_tree(tmp_path)
    monkeypatch.chdir(tree)
    _install_fake_clip()
    expected = load_expected()
    assert stage2.main(["--config", cfg_name, "--cls", cls]) == 0
    assert refine.main(["--config", cfg_name, "--cls", cls]) == 0
    for sid in ("scene0001_00", "scene0002_00", "scene0003_00"):
        got2 = torch.load(os.path.join(tree, out2, cls, f"{sid}.pth"), weights_only=False)
        gotf = torch.load(os.path.join(tree, outf, cls, f"{sid}.pth"), weights_only=False)
        check_against_expected(got2, expected[(cfg_name, cls, "mask_3d", sid)], f"stage2 {cls} {sid}")
        check_against_expected(gotf, expected[(cfg_name, cls, "final", sid)], f"final {cls} {sid}")
    assert os.path.exists(os.path.join(tree, "checkpoints", f"projection_2d_to_3d_checkpoint_{cls}.yaml"))


@pytest.mark.parametrize("cfg_name,out2", [("config.yaml", "output/mask_3d"), ("config_occ.yaml", "output_occ/mask_3d")])
def test_multi_query_run_equals_one_reference_run_per_class(cuda_dev, tmp_path, monkeypatch, cfg_name, out2):
    """Two text queries lifted in ONE pass per scene (shared projection / visibility / viewed counts, per-query
    votes) must write what the reference wrote when it was run once per class."""
    from bff_b200 import stage2
    tree = copy_tree(tmp_path)
    monkeypatch.chdir(tree)
    expected = load_expected()
    classes = ["office chair", "table"]
    assert stage2.main_multi(["--config", cfg_name, "--cls"] + classes) == 0
    for cls in classes:
        for sid in ("scene0001_00", "scene0002_00", "scene0003_00"):
            got = torch.load(os.path.join(tree, out2, cls, f"{sid}.pth"), weights_only=False)
            check_against_expected(got, expected[(cfg_name, cls, "mask_3d", sid)], f"multi stage2 {cls} {sid}")


@pytest.mark.parametrize("in_kernel", [True, False])
def test_cli_with_depth_upsampling_matches_reference(cuda_dev, tmp_path, monkeypatch, in_kernel):
    """The reference's default situation (configs/config.yaml:19-20): depth PNGs smaller than the masks, every frame
    resized by cv2.resize (:436, :563). tests/golden/tree_up + e2e_up.pth come from the unmodified reference CLIs
    (oracle/make_golden_up.py). in_kernel=True: the lift kernel samples the native uint16 depth itself
    (BFF_DEPTH_U16_BILINEAR, gated at run time against this machine's cv2.resize); False: cv2.resize on the host."""
    from bff_b200 import refine, scene_io, stage2
    tree = copy_tree(tmp_path, "tree_up")
    monkeypatch.chdir(tree)
    monkeypatch.setattr(scene_io, "IN_KERNEL_RESIZE", in_kernel)
    _install_fake_clip()
    expected = load_expected("e2e_up.pth")
    cls = "office chair"
    assert stage2.main(["--config", "config.yaml", "--cls", cls]) == 0
    assert refine.main(["--config", "config.yaml", "--cls", cls]) == 0
    if in_kernel:
        assert scene_io._resize_checked and all(scene_io._resize_checked.values()), \
            "the in-kernel resize disagrees with this machine's cv2.resize: the fused depth path was not exercised"
    for sid in ("scene0001_00", "scene0002_00"):
        got2 = torch.load(os.path.join(tree, "output/mask_3d", cls, f"{sid}.pth"), weights_only=False)
        gotf = torch.load(os.path.join(tree, "output/final_output", cls, f"{sid}.pth"), weights_only=False)
        check_against_expected(got2, expected[("config.yaml", cls, "mask_3d", sid)], f"stage2 {sid}")
        check_against_expected(gotf, expected[("config.yaml", cls, "final", sid)], f"final {sid}")


@pytest.mark.parametrize("mode", ["scene", "frame"])
def test_two_rank_cli_matches_reference(cuda_dev, tmp_path, mode):
    """torchrun, 2 ranks over NCCL, the drop-in lifting CLI on the golden tree: whole scenes per rank (LPT, config 5) and
    frames sharded per scene (vote all-reduce, row all-gather, all-pairs IoU split over the ranks) must both write the
    reference's files."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    from conftest import ROOT
    tree = copy_tree(tmp_path)
    expected = load_expected()
    cls = "office chair"
    env = dict(os.environ, BFF_DIST_MODE=mode, BFF_PAIRS_MIN_ROWS="2", PYTHONPATH=os.path.join(ROOT, "beyond-fixed-forms_b200"))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29571" if mode == "scene" else "29572", os.path.join(ROOT, "tools", "projection_2d_to_3d.py"),
           "--config", "config.yaml", "--cls", cls]
    r = subprocess.run(cmd, cwd=tree, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    for sid in ("scene0001_00", "scene0002_00", "scene0003_00"):
        got2 = torch.load(os.path.join(tree, "output/mask_3d", cls, f"{sid}.pth"), weights_only=False)
        check_against_expected(got2, expected[("config.yaml", cls, "mask_3d", sid)], f"{mode} stage2 {sid}")
    import yaml
    ck = yaml.safe_load(open(os.path.join(tree, "checkpoints", f"projection_2d_to_3d_checkpoint_{cls}.yaml")))
    assert ck and all(v is True for v in ck.values())


@pytest.mark.parametrize("cfg_name,out2,outf", [("config.yaml", "output/mask_3d", "output/final_output"),
                                                 ("config_occ.yaml", "output_occ/mask_3d", "output_occ/final_output")])
def test_multi_class_driver_equals_one_reference_run_per_class(cuda_dev, tmp_path, monkeypatch, cfg_name, out2, outf):
    """tools/run_evl_b200.py (bff_b200/run_classes.py): lifting + refinement for all classes of the tree in one process --
    shared projection per scene, stage-1 masks decoded once per scene, one similarity cache -- must write the files the
    unmodified reference wrote with one process per class and stage, and the orchestrator's checkpoint."""
    import yaml
    from bff_b200 import run_classes
    tree = copy_tree(tmp_path)
    monkeypatch.chdir(tree)
    _install_fake_clip()
    expected = load_expected()
    classes = ["office chair", "table"]
    ck = run_classes.run(cfg_name, classes, batch=2)
    assert all(ck[c] == {"projection_2d_to_3d": True, "refinement": True} for c in classes)
    assert yaml.safe_load(open("process_checkpoint.txt")) == ck
    for cls in classes:
        for sid in ("scene0001_00", "scene0002_00", "scene0003_00"):
            got2 = torch.load(os.path.join(tree, out2, cls, f"{sid}.pth"), weights_only=False)
            gotf = torch.load(os.path.join(tree, outf, cls, f"{sid}.pth"), weights_only=False)
            check_against_expected(got2, expected[(cfg_name, cls, "mask_3d", sid)], f"driver stage2 {cls} {sid}")
            check_against_expected(gotf, expected[(cfg_name, cls, "final", sid)], f"driver final {cls} {sid}")
    # a second run finds everything done and touches nothing
    before = {p: os.path.getmtime(p) for p in (os.path.join(tree, outf, classes[0], "scene0001_00.pth"),)}
    run_classes.run(cfg_name, classes, batch=2)
    assert all(os.path.getmtime(p) == t for p, t in before.items())
