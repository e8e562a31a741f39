"""CPU: the dictionary-building half of bff_b200/evaluator.py (label encoding, ground-truth lists, table-driven matching)
against the goldens of the unmodified reference evaluator (tests/golden/eval.pth), with the intersection tables supplied
by plain numpy instead of the CUDA kernels -- the host logic is checked without a GPU, the kernels in test_evaluator_gpu.py."""
import os

import numpy as np
import torch

from conftest import GOLDEN
from test_evaluator_gpu import _equal, _evaluator, _get_instances


def _numpy_tables(pred_masks, gts, instance_ids, void, device):
    inside = [np.not_equal(m, 0) for m in pred_masks]
    inter = np.asarray([[np.count_nonzero((gts == i) & m) for i in instance_ids] for m in inside], dtype=np.int64).reshape(len(inside), len(instance_ids))
    return (inter, np.asarray([np.count_nonzero(void & m) for m in inside], dtype=np.int64),
            np.asarray([np.count_nonzero(m) for m in inside], dtype=np.int64))


def test_matching_dictionaries_equal_reference_with_numpy_tables():
    from bff_b200 import evaluator
    G = torch.load(os.path.join(GOLDEN, "eval.pth"), weights_only=False)
    ev = _evaluator(G["class_labels"])
    for scan in G["scans"]:
        gt2pred, pred2gt = evaluator.assign_instances_for_scan(ev, scan["preds"], scan["gts_sem"].copy(), scan["gts_ins"].copy(),
                                                               _get_instances, tables=_numpy_tables)
        _equal({k: v for k, v in gt2pred.items() if v}, scan["gt2pred"], "gt2pred")
        _equal({k: v for k, v in pred2gt.items() if v}, scan["pred2gt"], "pred2gt")


def test_class_agnostic_lists_and_label_encoding():
    from bff_b200 import evaluator
    G = torch.load(os.path.join(GOLDEN, "eval.pth"), weights_only=False)
    scan = G["scans"][0]
    ev = _evaluator(G["class_labels"])
    code = evaluator.encode_ground_truth(ev, scan["gts_sem"].copy(), scan["gts_ins"].copy())
    sem = scan["gts_sem"] - 2 + 1                       # scannetv2_inst_eval.py:268-286, spelled out
    sem[sem < 0] = 0
    want = sem * 1000 + (scan["gts_ins"] + 1)
    want[(scan["gts_ins"] + 1) < 0] = 0
    np.testing.assert_array_equal(code, want)
    assert code.dtype == want.dtype
    ev.use_label = False
    gt2pred, pred2gt = evaluator.assign_instances_for_scan(ev, scan["preds"], scan["gts_sem"].copy(), scan["gts_ins"].copy(),
                                                           _get_instances, tables=_numpy_tables)
    first = G["class_labels"][0]
    assert list(gt2pred.keys()) == [first] and all(p["label_id"] is None for p in pred2gt[first])
    assert [p["pred_id"] for p in pred2gt[first]] == list(range(len(pred2gt[first])))
    assert sum(len(v) for k, v in pred2gt.items() if k != first) == 0
    n_gt = sum(len(v) for v in _get_instances(code, ev.valid_class_ids, ev.valid_class_labels, ev.id2label).values())
    assert len(gt2pred[first]) == n_gt
    for p in pred2gt[first]:                            # every match appears on both sides with the same numbers
        for g in p["matched_gt"]:
            assert any(q["pred_id"] == p["pred_id"] and q["intersection"] == g["intersection"] for q in g["matched_pred"])
