"""Opt-in accelerated matching step of the ScanNet200 evaluator.

The reference's ScanNetEval.assign_instances_for_scan (evaluation/eval/scannetv2_inst_eval.py:264-361) computes, for
every prediction mask and every ground-truth instance of its class, `np.count_nonzero(np.logical_and(gts == id, mask))`
-- O(P * G * N) numpy passes per scene. Here the ground-truth instances and the void mask become packed rows once
(bff_labels_to_rows, bff_pack_bits) and ALL intersections and vertex counts are one bff_mask_iou call; the matching
dictionaries are then filled exactly like the reference fills them (same keys, order, Python float IoU). The evaluator
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


def assign_instances_for_scan(ev, preds, gts_sem, gts_ins, get_instances: Callable, device=None):
    """Drop-in body of ScanNetEval.assign_instances_for_scan (scannetv2_inst_eval.py:264-361). `ev` is the evaluator
    object (dataset_name, encode_value, valid_class_ids, valid_class_labels, id2label, eval_class_labels, use_label,
    min_region_sizes); `get_instances` is the reference's evaluation.eval.instance_eval_util.get_instances."""
    device = device or torch.device("cuda", torch.cuda.current_device())
    # :268-286 label encoding, exactly as the reference (numpy int arithmetic on copies)
    if ev.dataset_name in ("scannetv2", "scannet200"):
        gts_sem = gts_sem - 2 + 1
    elif ev.dataset_name in ("replica", "stpls3d"):
        gts_sem = gts_sem - 1 + 1
    elif ev.dataset_name == "scannetpp":
        gts_sem = gts_sem - 105 + 1
    else:
        gts_sem = gts_sem + 1
    gts_sem[gts_sem < 0] = 0
    gts_ins = gts_ins + 1
    ignore_inds = gts_ins < 0
    gts = gts_sem * ev.encode_value + gts_ins
    gts[ignore_inds] = 0
    gt_instances = get_instances(gts, ev.valid_class_ids, ev.valid_class_labels, ev.id2label, dataset=ev.dataset_name)
    if ev.use_label:
        gt2pred = deepcopy(gt_instances)
        for label in gt2pred:
            for gt in gt2pred[label]:
                gt["matched_pred"] = []
    else:
        gt2pred = {}
        agnostic = []
        for _, instances in gt_instances.items():
            agnostic += deepcopy(instances)
        for gt in agnostic:
            gt["matched_pred"] = []
        gt2pred[ev.eval_class_labels[0]] = agnostic
    pred2gt = {label: [] for label in ev.eval_class_labels}
    bool_void = np.logical_not(np.isin(gts // ev.encode_value, ev.valid_class_ids))           # :312 (np.in1d)
    # every distinct ground-truth instance id gets one packed row; all intersections in one launch
    all_ids, col = [], {}
    for label in gt2pred:
        for gt in gt2pred[label]:
            if gt["instance_id"] not in col:
                col[gt["instance_id"]] = len(all_ids)
                all_ids.append(int(gt["instance_id"]))
    kept = []
    for pred in preds:
        if ev.use_label and pred["label_id"] not in ev.id2label:
            continue
        assert pred["pred_mask"].shape[0] == gts.shape[0]
        kept.append(pred)
    inter, void_inter, counts = intersection_tables([p["pred_mask"] for p in kept], gts, all_ids, bool_void, device)
    num_pred_instances = 0
    for k, pred in enumerate(kept):
        label_id = pred["label_id"]
        label_name = ev.id2label[label_id] if ev.use_label else ev.eval_class_labels[0]
        num = counts[k]                                   # numpy integer scalars like np.count_nonzero returns (:332)
        if num < ev.min_region_sizes[0]:
            continue
        pred_instance = {"filename": "{}_{}".format(pred["scan_id"], num_pred_instances), "pred_id": num_pred_instances,
                         "label_id": label_id if ev.use_label else None, "vert_count": num, "confidence": pred["conf"],
                         "void_intersection": void_inter[k]}
        matched_gt = []
        for gt_num, gt_inst in enumerate(gt2pred[label_name]):
            intersection = inter[k, col[gt_inst["instance_id"]]]
            if intersection > 0:
                gt_copy, pred_copy = gt_inst.copy(), pred_instance.copy()
                gt_copy["intersection"] = intersection
                pred_copy["intersection"] = intersection
                iou = float(intersection) / (gt_copy["vert_count"] + pred_copy["vert_count"] - intersection)
                gt_copy["iou"] = iou
                pred_copy["iou"] = iou
                matched_gt.append(gt_copy)
                gt2pred[label_name][gt_num]["matched_pred"].append(pred_copy)
        pred_instance["matched_gt"] = matched_gt
        num_pred_instances += 1
        pred2gt[label_name].append(pred_instance)
    return gt2pred, pred2gt


def install(scan_eval_cls, get_instances: Callable, device=None):
    """Swap the matching step on the reference's ScanNetEval class (evaluate() calls it per scene, :585)."""
    def method(self, preds, gts_sem, gts_ins):
        return assign_instances_for_scan(self, preds, gts_sem, gts_ins, get_instances, device)
    scan_eval_cls.assign_instances_for_scan = method
    return scan_eval_cls
