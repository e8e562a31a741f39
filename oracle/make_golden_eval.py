"""TEST INFRASTRUCTURE ONLY -- golden vectors for the evaluator's matching step from the UNMODIFIED reference
(evaluation/eval/scannetv2_inst_eval.py:264-361, ScanNetEval.assign_instances_for_scan, imported from /root/reference
with the stubs of oracle/run_reference.py plus an empty `plyfile`). Writes tests/golden/eval.pth:
inputs (pred dicts, gts_sem, gts_ins per scan) and the (gt2pred, pred2gt) the reference returned.
Run in the build container: python oracle/make_golden_eval.py"""
from __future__ import annotations

import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import run_reference as RR  # noqa: E402


def import_evaluator():
    with RR.reference_env():
        if "plyfile" not in sys.modules:
            m = types.ModuleType("plyfile")
            m.PlyData = object
            sys.modules["plyfile"] = m
        E = importlib.import_module("evaluation.eval.scannetv2_inst_eval")
        U = importlib.import_module("evaluation.eval.instance_eval_util")
        D = importlib.import_module("evaluation.dataset.scannet200")
    return E, U, D


def make_scan(seed, n, n_inst, n_pred, n_classes):
    """Contiguous ground-truth instances (raw semantic ids 2.., instance ids 0.., -1 = ignore) and predictions that overlap
    them partly, plus small (< 100 points), empty and unknown-label predictions."""
    rng = np.random.default_rng(seed)
    cuts = np.sort(rng.choice(np.arange(50, n - 50), n_inst - 1, replace=False))
    seg = np.searchsorted(cuts, np.arange(n), side="right")
    cls = rng.integers(0, n_classes, n_inst)
    gts_sem = (cls[seg] + 2).astype(np.int32)
    gts_ins = seg.astype(np.int32)
    hole = rng.random(n) < 0.03
    gts_ins[hole] = -2                                   # ignored points (:279-286)
    gts_sem[rng.random(n) < 0.02] = 0                    # unlabeled -> void
    preds = []
    for k in range(n_pred):
        g = int(rng.integers(0, n_inst))
        m = (seg == g) & (rng.random(n) < 0.8)
        m |= rng.random(n) < 0.01
        if k % 7 == 3:
            m = np.zeros(n, bool)
            m[:int(rng.integers(1, 60))] = True          # below min_region_size
        label = int(cls[g]) if k % 5 else int(rng.integers(0, n_classes))
        if k % 11 == 10:
            label = 5000                                 # not in id2label -> skipped
        preds.append({"scan_id": f"scene{seed:04d}_00", "label_id": float(label) + 1, "conf": 1.0, "pred_mask": m.astype(np.uint8)})
    return preds, gts_sem, gts_ins


def main():
    E, U, D = import_evaluator()
    ev = E.ScanNetEval(class_labels=D.INSTANCE_CAT_SCANNET_200, dataset_name="scannet200")
    scans = []
    for seed, n, n_inst, n_pred in [(1, 6000, 14, 20), (2, 4500, 9, 12), (3, 5003, 11, 0)]:
        preds, sem, ins = make_scan(seed, n, n_inst, n_pred, 30)
        gt2pred, pred2gt = ev.assign_instances_for_scan(preds, sem.copy(), ins.copy())
        # keep only classes that have something, the dictionaries list all 198 otherwise
        scans.append({"preds": preds, "gts_sem": sem, "gts_ins": ins,
                      "gt2pred": {k: v for k, v in gt2pred.items() if v}, "pred2gt": {k: v for k, v in pred2gt.items() if v}})
        print(seed, sum(len(v) for v in gt2pred.values()), "gt instances,", sum(len(v) for v in pred2gt.values()), "kept predictions,",
              sum(len(p["matched_gt"]) for v in pred2gt.values() for p in v), "matches")
    torch.save({"scans": scans, "class_labels": list(D.INSTANCE_CAT_SCANNET_200)}, os.path.join(ROOT, "tests", "golden", "eval.pth"))


if __name__ == "__main__":
    main()
