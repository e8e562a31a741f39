"""Opt-in accelerated matching step of the ScanNet200 evaluator.

The reference's ScanNetEval.assign_instances_for_scan (evaluation/eval/scannetv2_inst_eval.py:264-361) computes, for
every prediction mask and every ground-truth instance of its class, `np.count_nonzero(np.logical_and(gts == id, mask))`
-- O(P * G * N) numpy passes per scene. Here the ground-truth instances and the void mask become packed rows once
(bff_labels_to_rows, bff_pack_bits) and ALL intersections and vertex counts are one bff_mask_iou call; the matching
dictionaries are then built from that table, visiting only the overlapping (prediction, instance) pairs, with the keys,
order and value types the reference produces (the goldens of the unmodified evaluator pin that). The evaluator
keeps consuming the same .pth files; `install(ScanNetEval)` swaps the method on the reference's own class, everything
after it (evaluate_matches, AP integration) stays the reference's code.

Integer results are exact; iou = float(I) / (gt_count + pred_count - I) is computed on the host in Python floats like
:353, so the dictionaries compare equal to the reference's."""
from __future__ import annotations

from copy import deepcopy
from typing import Callable, Sequence

import numpy as np
import torch

from . import ops


def intersection_tables(pred_masks: Sequence[np.ndarray], gts: np.ndarray, instance_ids: Sequence[int], void: np.ndarray, device):
    """pred_masks: P arrays (N,) (non-zero = inside, :331); gts (N,) encoded labels; instance_ids (G,); void (N,) bool.
    Returns numpy int64 (P, G) intersections, (P,) void intersections, (P,) vertex counts."""
    p, n = len(pred_masks), int(gts.shape[0])
    if p == 0:
        return np.zeros((0, len(instance_ids)), np.int64), np.zeros(0, np.int64), np.zeros(0, np.int64)
    dense = torch.from_numpy(np.ascontiguousarray(np.stack([np.not_equal(m, 0) for m in pred_masks]).astype(np.uint8))).to(device)
    pred_rows = ops.pack_bits(dense)
    ids = torch.as_tensor(list(instance_ids) if len(instance_ids) else [0], dtype=torch.int64, device=device)
    gt_rows = ops.labels_to_rows(torch.from_numpy(np.ascontiguousarray(gts, dtype=np.int64)).to(device), ids)
    void_row = ops.pack_bits(torch.from_numpy(np.ascontiguousarray(void, dtype=np.uint8)[None]).to(device))
    inter, pop_pred, _, _ = ops.mask_iou(pred_rows, torch.cat([gt_rows, void_row]), want_iou=False)
    inter = inter.cpu().numpy().astype(np.int64)
    g = len(instance_ids)
    return inter[:, :g], inter[:, -1], pop_pred.cpu().numpy().astype(np.int64)


# semantic-label offset per dataset before the evaluator's "+1" (:268-280); unknown names use 0
_SEMANTIC_OFFSET = {"scannetv2": 2, "scannet200": 2, "replica": 1, "stpls3d": 1, "scannetpp": 105}


def encode_ground_truth(ev, gts_sem: np.ndarray, gts_ins: np.ndarray) -> np.ndarray:
    """The evaluator's per-vertex code `semantic * encode_value + instance` (:268-286): semantic ids shifted to start at
    1 with negatives clamped to 0, instance ids shifted by one, vertices without an instance coded 0. Integer arithmetic,
    so the single shift below equals the reference's two-step one."""
    sem = gts_sem - (_SEMANTIC_OFFSET.get(ev.dataset_name, 0) - 1)
    sem = np.where(sem < 0, 0, sem).astype(sem.dtype, copy=False)
    ins = gts_ins + 1
    code = sem * ev.encode_value + ins
    code[ins < 0] = 0
    return code


def _ground_truth_lists(ev, gt_instances: dict) -> dict:
    """label -> list of ground-truth records (deep copies, each with an empty `matched_pred`); one list under the first
    evaluation label when the evaluation is class-agnostic (:290-305)."""
    if ev.use_label:
        lists = deepcopy(gt_instances)
    else:
        lists = {ev.eval_class_labels[0]: [rec for records in gt_instances.values() for rec in deepcopy(records)]}
    for records in lists.values():
        for rec in records:
            rec["matched_pred"] = []
    return lists


def assign_instances_for_scan(ev, preds, gts_sem, gts_ins, get_instances: Callable, device=None, tables: Callable = None):
    """What ScanNetEval.assign_instances_for_scan returns (scannetv2_inst_eval.py:264-361): (gt2pred, pred2gt) with the same
    keys, order and value types, built from ONE table of intersections instead of a numpy pass per (prediction, instance).
    `ev` is the evaluator object (dataset_name, encode_value, valid_class_ids, valid_class_labels, id2label,
    eval_class_labels, use_label, min_region_sizes); `get_instances` is the reference's
    evaluation.eval.instance_eval_util.get_instances; `tables` defaults to the CUDA `intersection_tables`."""
    tables = tables or intersection_tables
    if tables is intersection_tables:
        device = device or torch.device("cuda", torch.cuda.current_device())
    gts = encode_ground_truth(ev, gts_sem, gts_ins)
    gt2pred = _ground_truth_lists(ev, get_instances(gts, ev.valid_class_ids, ev.valid_class_labels, ev.id2label, dataset=ev.dataset_name))
    pred2gt = {label: [] for label in ev.eval_class_labels}
    void = ~np.isin(gts // ev.encode_value, ev.valid_class_ids)                    # :312
    # table columns: the distinct ground-truth instance ids; per label the columns of its records, in list order
    column: dict = {}
    label_cols = {label: np.asarray([column.setdefault(int(rec["instance_id"]), len(column)) for rec in records], dtype=np.int64)
                  for label, records in gt2pred.items()}
    scored = [p for p in preds if not ev.use_label or p["label_id"] in ev.id2label]
    for p in scored:
        assert p["pred_mask"].shape[0] == gts.shape[0]
    inter, void_inter, vert_count = tables([p["pred_mask"] for p in scored], gts, list(column), void, device)
    agnostic = ev.eval_class_labels[0]
    for k in np.flatnonzero(vert_count >= ev.min_region_sizes[0]):                # :333 drops the smaller predictions
        pred = scored[k]
        label = ev.id2label[pred["label_id"]] if ev.use_label else agnostic
        pred_id = sum(len(v) for v in pred2gt.values())
        record = {"filename": "{}_{}".format(pred["scan_id"], pred_id), "pred_id": pred_id,
                  "label_id": pred["label_id"] if ev.use_label else None, "vert_count": vert_count[k],
                  "confidence": pred["conf"], "void_intersection": void_inter[k]}
        matched = []
        row = inter[k, label_cols[label]]
        for j in np.flatnonzero(row > 0):                                         # only the overlapping instances are visited
            both = row[j]
            gt_rec = gt2pred[label][j]
            pair = {"intersection": both, "iou": float(both) / (gt_rec["vert_count"] + record["vert_count"] - both)}
            matched.append({**gt_rec, **pair})
            gt_rec["matched_pred"].append({**record, **pair})
        record["matched_gt"] = matched
        pred2gt[label].append(record)
    return gt2pred, pred2gt


def install(scan_eval_cls, get_instances: Callable, device=None):
    """Swap the matching step on the reference's ScanNetEval class (evaluate() calls it per scene, :585)."""
    def method(self, preds, gts_sem, gts_ins):
        return assign_instances_for_scan(self, preds, gts_sem, gts_ins, get_instances, device)
    scan_eval_cls.assign_instances_for_scan = method
    return scan_eval_cls
