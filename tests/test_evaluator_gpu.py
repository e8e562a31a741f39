"""GPU: the accelerated matching step of the evaluator (bff_b200/evaluator.py) fills gt2pred / pred2gt
exactly like the unmodified reference's ScanNetEval.assign_instances_for_scan did on the same inputs
(tests/golden/eval.pth, oracle/make_golden_eval.py): same keys, order, integer counts and Python-float IoUs."""
import os
import types

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _get_instances(ids, class_ids, class_labels, id2label, coords=None, dataset="scannet200"):
    """evaluation/eval/instance_eval_util.py:159-174 with Instance.to_dict (:71-111) inlined -- an INPUT of the matching
    step (the product takes the reference's own function as an argument)."""
    instances = {label: [] for label in class_labels}
    for i in np.unique(ids):
        if i == 0:
            continue
        label_id = int(i // 1000)
        if label_id in class_ids:
            instances[id2label[label_id]].append({"instance_id": int(i), "label_id": label_id, "vert_count": int((ids == i).sum()),
                                                  "med_dist": -1, "dist_conf": 0.0, "box": np.zeros(6)})
    return instances


def _evaluator(class_labels):
    """The fields ScanNetEval.__init__ sets (scannetv2_inst_eval.py:22-54) for dataset_name='scannet200'."""
    ids = np.arange(len(class_labels)) + 1
    return types.SimpleNamespace(dataset_name="scannet200", encode_value=1000, valid_class_labels=class_labels, valid_class_ids=ids,
                                 id2label={ids[i]: class_labels[i] for i in range(len(ids))}, eval_class_labels=class_labels,
                                 use_label=True, min_region_sizes=np.array([100]))


def _equal(a, b, path=""):
    if isinstance(a, dict):
        assert isinstance(b, dict) and list(a.keys()) == list(b.keys()), path
        for k in a:
            _equal(a[k], b[k], f"{path}/{k}")
    elif isinstance(a, (list, tuple)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _equal(x, y, f"{path}[{i}]")
    elif isinstance(a, np.ndarray):
        np.testing.assert_array_equal(a, b, err_msg=path)
    else:
        assert a == b and type(a) is type(b), (path, a, b)


def test_matching_tables_equal_reference(cuda_dev):
    from bff_b200 import evaluator
    G = torch.load(os.path.join(GOLDEN, "eval.pth"), weights_only=False)
    ev = _evaluator(G["class_labels"])
    n_matches = 0
    for scan in G["scans"]:
        gt2pred, pred2gt = evaluator.assign_instances_for_scan(ev, scan["preds"], scan["gts_sem"].copy(), scan["gts_ins"].copy(),
                                                               _get_instances, cuda_dev)
        assert set(gt2pred.keys()) == set(G["class_labels"]) == set(pred2gt.keys())
        _equal({k: v for k, v in gt2pred.items() if v}, scan["gt2pred"], "gt2pred")
        _equal({k: v for k, v in pred2gt.items() if v}, scan["pred2gt"], "pred2gt")
        n_matches += sum(len(p["matched_gt"]) for v in pred2gt.values() for p in v)
    assert n_matches >= 25


def test_labels_to_rows_and_tables_against_numpy(cuda_dev):
    from bff_b200 import evaluator, ops
    from helpers import unpack_rows
    rng = np.random.default_rng(9)
    n = 70_001
    gts = rng.integers(0, 1500, n).astype(np.int64) * 1000 + rng.integers(0, 3, n)
    ids = np.unique(gts)[:1300]                                              # > 1024: two shared-memory passes
    rows = ops.labels_to_rows(torch.from_numpy(gts).to(cuda_dev), torch.from_numpy(ids).to(cuda_dev))
    got = unpack_rows(rows.cpu().numpy(), n)
    for g in (0, 1, 640, 1023, 1024, 1299):
        np.testing.assert_array_equal(got[g], gts == ids[g])
    preds = [(rng.random(n) < p).astype(np.uint8) * 3 for p in (0.5, 0.01, 0.0)]
    void = rng.random(n) < 0.1
    inter, vint, cnt = evaluator.intersection_tables(preds, gts, ids[:40].tolist(), void, cuda_dev)
    for k, m in enumerate(preds):
        assert cnt[k] == np.count_nonzero(m) and vint[k] == np.count_nonzero(void & (m != 0))
        for g in range(40):
            assert inter[k, g] == np.count_nonzero((gts == ids[g]) & (m != 0))
