"""CPU: the numpy/torch oracle port (oracle/reference_port.py), driven over the committed input tree,
reproduces what the UNMODIFIED reference CLIs wrote for that tree (tests/golden/e2e.pth)."""
import os

import pytest
import yaml

from helpers import check_against_expected, copy_tree, load_expected, port_run_class


@pytest.mark.parametrize("cfg_name", ["config.yaml", "config_occ.yaml"])
@pytest.mark.parametrize("cls", ["office chair", "table"])
def test_port_reproduces_reference_outputs(tmp_path, monkeypatch, cfg_name, cls):
    tree = copy_tree(tmp_path)
    monkeypatch.chdir(tree)
    cfg = yaml.safe_load(open(os.path.join(tree, cfg_name)))
    expected = load_expected()
    stage2, finals = port_run_class(cfg, cls)
    n = 0
    for sid in stage2:
        check_against_expected(stage2[sid], expected[(cfg_name, cls, "mask_3d", sid)], f"stage2 {cls} {sid}")
        check_against_expected(finals[sid], expected[(cfg_name, cls, "final", sid)], f"final {cls} {sid}")
        n += 1
    assert n == 3
