"""Stage-1 / stage-2 refinement on packed point masks: the body of tools/refinement.py
(:135-428). Mask algebra (RLE ingest, IoU, OR-merges) runs in libbff.so; matching decisions are
taken on the small IoU matrices with the same torch CPU ops as the reference."""
from __future__ import annotations

import os
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _log, ops, scene_io
from ._log import log
from .rle import pairs_for_device

_HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(_HERE, "stage1_classes.txt")) as _fh:
    # index -> label table of the stage-1 (Open3DIS) results, refinement.py:58-62
    STAGE1_CLASS_NAMES: List[str] = [l.strip() for l in _fh if l.strip()]


MAX_DEVICE_ROWS, MAX_FUSED_ROWS = 40960, 65535   # stage-2 mask counts bff_refine_device / bff_refine_match accept (include/bff.h);
                                                 # larger jobs take the next path down, the step-by-step one has no limit
FUSED = True     # bff_refine_match (C++ decisions next to the launches); False = one Python/torch call per step
DEVICE = True    # bff_refine_device (every decision a kernel, one synchronisation per scene); takes precedence over FUSED


def idx_to_label(idx) -> str:
    return STAGE1_CLASS_NAMES[idx]


def iou_between_stages(rows1: torch.Tensor, rows2: torch.Tensor) -> torch.Tensor:
    """calculate_iou_between_stages (refinement.py:69-90): (m, n) float32 CPU matrix."""
    _, _, _, iou = ops.mask_iou(rows1, rows2)
    return iou.cpu().T


def stage1_rows(stage1: dict, device) -> torch.Tensor:
    """RLE stage-1 masks -> packed rows on the device (refinement.py:26-39,186-189 without the
    dense uint8 intermediate)."""
    rles = list(stage1["ins"])
    if not rles:
        return torch.zeros((0, 0), dtype=torch.int32, device=device)
    counts, pair_off, length = pairs_for_device(rles)
    return ops.rle_pairs_to_bits_dev(torch.from_numpy(counts).to(device), torch.from_numpy(pair_off).to(device), length)


def pass1_scene(stage1: dict, stage2: dict, query: str, stage1_iou_thres: float,
                sim_fn: Callable[[str, str], float], device, s1_rows: Optional[torch.Tensor] = None) -> dict:
    """Pass 1 for one scene (refinement.py:182-312). Returns packed rows + the small vectors pass 2 needs.
    `s1_rows`: the stage-1 masks already decoded to packed rows on the device (modified in place)."""
    q_ = query.replace(" ", "_")
    s1 = stage1_rows(stage1, device) if s1_rows is None else s1_rows
    n_points = int(stage1["ins"][0]["length"]) if len(stage1["ins"]) else 0
    labels = [STAGE1_CLASS_NAMES[i] for i in torch.as_tensor(stage1["final_class"]).tolist()]
    if len(stage2["conf"]) == 0:
        log("Empty stage 2 mask")
        other = [i for i, l in enumerate(labels) if l == q_]
        return {"ious": [], "sims": [], "m1": None, "m2": None, "conf2": [], "n_points": n_points,
                "other": s1[ops.small_to_device(torch.tensor(other, dtype=torch.long), device)] if other else s1[:0]}
    s2 = stage2["rows"] if "rows" in stage2 else ops.pack_bits(stage2["ins"].to(device))
    s2_conf = stage2["conf"]
    if DEVICE and isinstance(s2_conf, torch.Tensor) and s2_conf.dtype == torch.float32 and s1.shape[0] > 0 and s2.shape[0] <= MAX_DEVICE_ROWS:
        return _pass1_device(s1, s2, s2_conf, labels, query, q_, stage1_iou_thres, sim_fn, n_points, device)
    if FUSED and s2_conf.dtype == torch.float32 and s2.shape[0] <= MAX_FUSED_ROWS:
        return _pass1_fused(s1, s2, s2_conf, labels, query, q_, stage1_iou_thres, sim_fn, n_points, device)
    mm = torch.argmax(iou_between_stages(s1, s2), dim=1)                       # :208-213
    sel = s1[mm.to(device)]
    g = iou_between_stages(sel, sel)                                           # :217-219
    g[range(len(mm)), range(len(mm))] = 0
    g = (g > stage1_iou_thres).to(int)
    best: list = []
    redirect = torch.ones(len(g), dtype=torch.int) * -1
    or_prog = []                                                               # sequential in-place ORs (:248)
    for i in range(len(g)):
        if redirect[i] != -1:
            best.append(mm[redirect[i]])
            continue
        best.append(mm[i])
        if g[i].sum() > 0:
            for j in torch.nonzero(g[i] == 1).reshape(-1).tolist():
                redirect[j] = i
                or_prog.append((int(mm[i]), int(mm[j])))
    ops.rows_or_into(s1, or_prog)
    log("DEBUG best_match_after_iou_check", best)
    best = torch.tensor(best)
    uniq, cnt = torch.unique(best, return_counts=True)                        # :258
    for u, c in zip(uniq, cnt):
        if c > 1:
            hit = best == u
            members = torch.nonzero(hit).reshape(-1).tolist()
            merged = ops.rows_or_reduce(s2, [members])
            mconf = s2_conf[hit].mean()
            keep = torch.nonzero(~hit).reshape(-1).to(device)
            s2 = torch.cat([s2[keep], merged])
            s2_conf = torch.cat([s2_conf[~hit], mconf.unsqueeze(0)])
            best = torch.cat([best[~hit], u.unsqueeze(0)])
    iou = iou_between_stages(s1, s2)                                           # :285-290
    mm = torch.argmax(iou, dim=1)
    other = [i for i, l in enumerate(labels) if l == q_ and i not in mm]      # :293
    matched_labels = [labels[k] for k in mm]
    sims = [float(sim_fn(query, l)) for l in matched_labels]                   # :297-302
    ious = iou[range(len(mm)), mm]
    log("iou for matched labels", ious)
    log("Matched labels:", matched_labels)
    log("Clip similarities:", sims)
    return {"ious": ious, "sims": sims, "m1": s1[mm.to(device)], "m2": s2, "conf2": s2_conf, "n_points": n_points,
            "other": s1[ops.small_to_device(torch.tensor(other, dtype=torch.long), device)] if other else s1[:0]}


class _RefineJob:
    """Descriptor and host output buffers of one bff_refine_match job."""

    def __init__(self, s1, s2, s2_conf, n_points, stage1_iou_thres):
        from . import _cabi
        n1, n2 = s1.shape[0], s2.shape[0]
        self.s1, self.s2, self.n_points = s1, s2, n_points
        self.conf32 = np.ascontiguousarray(s2_conf.numpy(), dtype=np.float32)
        self.s2_out = torch.empty_like(s2)
        self.n2_out = np.zeros(1, np.int32)
        self.conf_out, self.match_iou = np.zeros(n2, np.float32), np.zeros(n2, np.float32)
        self.match, self.best_log = np.zeros(n2, np.int32), np.zeros(n2, np.int32)
        d = self.desc = _cabi.RefineDesc()
        d.s1_rows, d.n1, d.s2_rows, d.n2, d.n_points = _cabi.ptr(s1), n1, _cabi.ptr(s2), n2, int(n_points)
        d.s2_conf, d.stage1_iou_thres = _cabi.ptr(self.conf32), float(np.float32(stage1_iou_thres))
        d.s2_out, d.n2_out, d.s2_conf_out = _cabi.ptr(self.s2_out), _cabi.ptr(self.n2_out), _cabi.ptr(self.conf_out)
        d.match, d.match_iou, d.best_match_log = _cabi.ptr(self.match), _cabi.ptr(self.match_iou), _cabi.ptr(self.best_log)

    def result(self, labels, query, q_, sim_fn, device) -> dict:
        from . import _cabi
        _cabi.launch_count += 14
        s1 = self.s1
        m = int(self.n2_out[0])
        mm = self.match[:m]
        if _log.VERBOSE:                     # the reference prints a list of 0-d tensors; do not build them for nobody
            log("DEBUG best_match_after_iou_check", [torch.tensor(int(b)) for b in self.best_log])
        matched = set(int(k) for k in mm)
        other = [i for i, l in enumerate(labels) if l == q_ and i not in matched]      # :293
        matched_labels = [labels[int(k)] for k in mm]
        sims = [float(sim_fn(query, l)) for l in matched_labels]                       # :297-302
        ious = torch.from_numpy(self.match_iou[:m].copy())
        log("iou for matched labels", ious)
        log("Matched labels:", matched_labels)
        log("Clip similarities:", sims)
        # matched rows and the class's other rows: one index upload, one gather
        pick = ops.small_to_device(torch.from_numpy(np.concatenate([mm.astype(np.int64), np.asarray(other, dtype=np.int64)])), device)
        sel = s1[pick]
        return {"ious": ious, "sims": sims, "m1": sel[:m], "m2": self.s2_out[:m], "conf2": torch.from_numpy(self.conf_out[:m].copy()),
                "n_points": self.n_points, "other": sel[m:]}


def _pass1_device(s1, s2, s2_conf, labels, query, q_, stage1_iou_thres, sim_fn, n_points, device) -> dict:
    """The matching part of pass 1 (refinement.py:208-312) with every decision on the device: bff_refine_device in
    select_mode 0, ONE synchronisation (the result block), then the similarity function for the MATCHED labels only, like
    the reference (:297-302; with CLIP every call is two text encodings). Pass 2 later runs bff_refine_select_device on the
    state left on the device (`rf`)."""
    from . import devpipe
    isq = ops.small_to_device(torch.tensor([l == q_ for l in labels], dtype=torch.uint8), device)
    sim0 = torch.zeros((len(labels),), dtype=torch.float64, device=device)           # filled in for the matches below
    cfg = {"stage1_iou_thres": stage1_iou_thres, "refinment_sim_percentile": 0.0, "refiment_iou_thres": 0.0}
    rf = devpipe.refine_device(s1, s2, int(s2.shape[0]), ops.small_to_device(s2_conf, device), isq, sim0, n_points, cfg, query, select_mode=0)
    h = rf.header()
    if _log.VERBOSE:
        log("DEBUG best_match_after_iou_check", [torch.tensor(int(b)) for b in h["best"]])
    matched_labels = [labels[int(k)] for k in h["match"]]
    sims = [float(sim_fn(query, l)) for l in matched_labels]                           # :297-302
    rf.set_sims(sims)
    ious = torch.from_numpy(h["match_iou"].copy())
    log("iou for matched labels", ious)
    log("Matched labels:", matched_labels)
    log("Clip similarities:", sims)
    return {"ious": ious, "sims": sims, "rf": rf, "conf2": torch.from_numpy(h["conf_out"].copy()), "n_points": n_points}


def _pass1_fused(s1, s2, s2_conf, labels, query, q_, stage1_iou_thres, sim_fn, n_points, device) -> dict:
    """The matching part of pass 1 (refinement.py:208-290) in one call: bff_refine_match."""
    from . import _cabi
    job = _RefineJob(s1, s2, s2_conf, n_points, stage1_iou_thres)
    _cabi.check(_cabi.lib().bff_refine_match(_cabi.ctx(s1.device.index or 0), job.desc, _cabi.current_stream()), "bff_refine_match")
    return job.result(labels, query, q_, sim_fn, device)


def pass1_scene_batch(stage1: dict, stage2_list: Sequence[dict], queries: Sequence[str], stage1_iou_thres: float,
                      sim_fn: Callable[[str, str], float], device, s1_rows_list: Sequence[torch.Tensor]) -> list:
    """pass1_scene for several queries of ONE scene: the fused jobs go through one bff_refine_match_batch call and share
    its synchronisations. `s1_rows_list[k]` is query k's OWN copy of the decoded stage-1 rows (pass 1 ORs into them,
    refinement.py:248; every reference process starts from the file). Returns one pass1_scene dict per query."""
    from . import _cabi
    n_points = int(stage1["ins"][0]["length"]) if len(stage1["ins"]) else 0
    labels = [STAGE1_CLASS_NAMES[i] for i in torch.as_tensor(stage1["final_class"]).tolist()]
    out = [None] * len(queries)
    at, jobs = [], []
    for k, (stage2, query) in enumerate(zip(stage2_list, queries)):
        fused = (FUSED and len(stage2["conf"]) and isinstance(stage2["conf"], torch.Tensor) and stage2["conf"].dtype == torch.float32
                 and len(stage2["conf"]) <= MAX_FUSED_ROWS)
        if not fused:
            out[k] = pass1_scene(stage1, stage2, query, stage1_iou_thres, sim_fn, device, s1_rows=s1_rows_list[k])
            continue
        s2 = stage2["rows"] if "rows" in stage2 else ops.pack_bits(stage2["ins"].to(device))
        at.append(k)
        jobs.append(_RefineJob(s1_rows_list[k], s2, stage2["conf"], n_points, stage1_iou_thres))
    if jobs:
        descs = (_cabi.RefineDesc * len(jobs))(*[j.desc for j in jobs])
        _cabi.check(_cabi.lib().bff_refine_match_batch(_cabi.ctx(jobs[0].s1.device.index or 0), descs, len(jobs), _cabi.current_stream()),
                    "bff_refine_match_batch")
        for k, j in zip(at, jobs):
            out[k] = j.result(labels, queries[k], queries[k].replace(" ", "_"), sim_fn, device)
    return out


def similarity_threshold(all_sims: Sequence[Sequence[float]], percentile: float) -> float:
    """Global over the scenes of the class (refinement.py:317-324). Raises IndexError like the
    reference when no scene produced a match."""
    u = sorted(set(s for sims in all_sims for s in sims))
    log("Unique similarities:", u)
    return u[int(len(u) * percentile)]


def pass2_scene(p1: dict, query: str, iou_thres: float, sim_thres: float, to_host: bool = True) -> dict:
    """Selection (refinement.py:330-426) -> the dict that is saved: the other stage-1 masks of the class first
    (conf 0.5), then per matched pair the stage-1 mask (IoU above the threshold and CLIP-similar enough), nothing
    (IoU above, not similar) or the stage-2 mask (IoU not above); confidences are the stage-2 ones."""
    if "rf" in p1:                               # state of bff_refine_device: the selection runs on the device too
        rf = p1["rf"]
        rf.desc.refine_iou_thres = float(np.float32(iou_thres))
        rf.select(sim_thres)
        return rf.final(to_host)
    out = {"ins": [], "conf": [], "final_class": []}
    n_other = p1["other"].shape[0]
    kept, from_s1 = [], []
    ious = p1["ious"]
    # float32 tensor element vs python float, like the reference (:360); one vectorised compare when the IoUs are a tensor
    above = (ious > iou_thres).tolist() if isinstance(ious, torch.Tensor) and ious.dim() == 1 else [bool(iou > iou_thres) for iou in ious]
    for k, hit in enumerate(above):
        if hit:
            if p1["sims"][k] < sim_thres:
                continue
            kept.append(k)
            from_s1.append(True)
        else:
            kept.append(k)
            from_s1.append(False)
    total = n_other + len(kept)
    if total == 0:
        return out
    out["final_class"] = [query] * total
    parts = [p1["other"]] if n_other else []
    if kept:
        m1, m2 = p1["m1"], p1["m2"]
        if len(kept) == m2.shape[0] and not any(from_s1):
            parts.append(m2)                     # every pair keeps its stage-2 mask: no index upload, no gather
        elif len(kept) == m1.shape[0] and all(from_s1):
            parts.append(m1)
        else:
            # one upload: row k of the result = row kept[k] of m1 (stage-1 mask) or of m2, taken from their concatenation
            pick = torch.tensor([k if f else k + m1.shape[0] for k, f in zip(kept, from_s1)], dtype=torch.long)
            parts.append(torch.cat([m1, m2])[ops.small_to_device(pick, m2.device)])
    rows = torch.cat(parts) if len(parts) > 1 else parts[0]
    ins = ops.unpack_bits(rows, p1["n_points"])
    out["ins"] = ops.to_host(ins) if to_host else ins
    conf2 = p1["conf2"]
    if isinstance(conf2, torch.Tensor) and conf2.dtype == torch.float32:
        out["conf"] = torch.cat([torch.full((n_other,), 0.5), conf2[torch.tensor(kept, dtype=torch.long)]])
    else:                                        # mixed dtypes: let torch.stack promote exactly like the reference (:412)
        out["conf"] = torch.stack([torch.tensor(0.5)] * n_other + [conf2[k] for k in kept])
    return out


class ClipTextSimilarity:
    """Cosine similarity of CLIP text embeddings (refinement.py:93-117), with a per-pair cache.
    CLIP is the reference's model and an INPUT to this path; it is used when importable."""

    def __init__(self, device):
        import clip  # the reference's dependency (OpenAI CLIP); out of scope to re-implement
        self._clip = clip
        self.device = device
        self.model, _ = clip.load("ViT-L/14", device=device)
        self._cache: Dict[tuple, float] = {}

    def __call__(self, query: str, label: str) -> float:
        key = (query, label)
        if key not in self._cache:
            a = self._clip.tokenize([query]).to(self.device)
            b = self._clip.tokenize([label]).to(self.device)
            with torch.no_grad():
                fa, fb = self.model.encode_text(a), self.model.encode_text(b)
            sim = fa @ fb.T
            sim = sim / (fa.norm(dim=-1, keepdim=True) * fb.norm(dim=-1, keepdim=True).T)
            self._cache[key] = sim.item()
        return self._cache[key]


def refine_class(cfg, cls: str, device, sim_fn: Callable[[str, str], float]):
    """Both passes over every scene of one class; writes <final_output_dir>/<cls>/<scene>.pth."""
    g = (lambda k: cfg[k]) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k))
    stage1_dir, mask_3d_dir = g("stage_1_results_dir"), g("mask_3d_dir")
    ckpt_path = f"./checkpoints/refinement_checkpoint_{cls}.yaml"             # :41-43
    ckpt = scene_io.read_checkpoint(ckpt_path)
    stage2_files = sorted(s for s in os.listdir(os.path.join(mask_3d_dir, cls)) if s.endswith("_00.pth"))
    per_scene = []
    for fn in stage2_files:
        scene_id = fn.replace(".pth", "")
        p1_path = os.path.join(stage1_dir, f"{scene_id}.pth")
        p2_path = os.path.join(mask_3d_dir, cls, f"{scene_id}.pth")
        if not (os.path.exists(p1_path) and os.path.exists(p2_path)):
            continue                                                           # :175-178 (B10: misaligns pass 2)
        per_scene.append(pass1_scene(scene_io.torch_load(p1_path), scene_io.torch_load(p2_path), cls,
                                     g("stage1_iou_thres"), sim_fn, device))
    sim_thres = similarity_threshold([p["sims"] for p in per_scene], g("refinment_sim_percentile"))
    for s, fn in enumerate(stage2_files):
        scene_id = fn.replace(".pth", "")
        p1 = per_scene[s]                                                      # positional, like the reference
        log("DEBUG all ious", p1["ious"])
        out = pass2_scene(p1, cls, g("refiment_iou_thres"), sim_thres)
        scene_io.save_result(out, g("final_output_dir"), cls, scene_id)
        if len(p1["ious"]) != 0 and len(out["final_class"]) != 0:
            ckpt[scene_id] = True                                              # :427-428
            scene_io.write_checkpoint(ckpt_path, ckpt)


def refine_classes(cfg, classes: Sequence[str], device, sim_fn: Callable[[str, str], float]):
    """refine_class for several classes in ONE process (the reference starts one process per class,
    evaluation/eval/run_evl.py:80, and each reloads CLIP, refinement.py:147): the scenes are walked once, every scene's
    stage-1 file is read and decoded once and handed to each class as a copy (pass 1 ORs into it, :248), and `sim_fn`
    (with its cache) is shared. Per class the outputs, the global threshold (:317-324) and the checkpoint file are
    exactly those of one refine_class call."""
    g = (lambda k: cfg[k]) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k))
    stage1_dir, mask_3d_dir = g("stage_1_results_dir"), g("mask_3d_dir")
    files = {cls: sorted(s for s in os.listdir(os.path.join(mask_3d_dir, cls)) if s.endswith("_00.pth")) for cls in classes}
    per_scene = {cls: [] for cls in classes}
    for fn in sorted(set(f for fs in files.values() for f in fs)):
        scene_id = fn.replace(".pth", "")
        p1_path = os.path.join(stage1_dir, f"{scene_id}.pth")
        stage1, master = None, None
        for cls in classes:
            if fn not in files[cls]:
                continue
            p2_path = os.path.join(mask_3d_dir, cls, fn)
            if not (os.path.exists(p1_path) and os.path.exists(p2_path)):
                continue                                                       # :175-178 (B10: misaligns pass 2)
            if stage1 is None:
                stage1 = scene_io.torch_load(p1_path)
                master = stage1_rows(stage1, device)
            log("Processing class:", cls.replace(" ", "_"), "scene", scene_id)
            per_scene[cls].append(pass1_scene(stage1, scene_io.torch_load(p2_path), cls, g("stage1_iou_thres"), sim_fn, device,
                                              s1_rows=master.clone()))
    for cls in classes:
        ckpt_path = f"./checkpoints/refinement_checkpoint_{cls}.yaml"
        ckpt = scene_io.read_checkpoint(ckpt_path)
        sim_thres = similarity_threshold([p["sims"] for p in per_scene[cls]], g("refinment_sim_percentile"))
        for s, fn in enumerate(files[cls]):
            scene_id = fn.replace(".pth", "")
            p1 = per_scene[cls][s]                                             # positional, like the reference
            log("DEBUG all ious", p1["ious"])
            out = pass2_scene(p1, cls, g("refiment_iou_thres"), sim_thres)
            scene_io.save_result(out, g("final_output_dir"), cls, scene_id)
            if len(p1["ious"]) != 0 and len(out["final_class"]) != 0:
                ckpt[scene_id] = True
                scene_io.write_checkpoint(ckpt_path, ckpt)


def main_multi(argv=None):
    """tools/refinement_multi.py: `--config C --cls q1 q2 ...` -- the outputs and checkpoints of one reference run per
    class, from one process."""
    import argparse
    import yaml
    parser = argparse.ArgumentParser(description="BeyondFF refinement for several classes at once")
    parser.add_argument("--config", type=str, required=True)
    parser.add_argument("--cls", type=str, nargs="+", required=True)
    args = parser.parse_args(argv)
    _log.VERBOSE = True
    cfg = yaml.safe_load(open(args.config, "r").read())
    if not torch.cuda.is_available():
        raise RuntimeError("bff_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    device = torch.device("cuda", 0)
    refine_classes(cfg, args.cls, device, ClipTextSimilarity(device))
    return 0


def main(argv=None):
    """CLI body of tools/refinement.py: same arguments, config keys and outputs."""
    import argparse
    import yaml
    parser = argparse.ArgumentParser(description="Configuration Beyond Fixed Forms")
    parser.add_argument("--config", type=str, required=True, help="Config")
    parser.add_argument("--cls", type=str, required=True, help="Class")
    args = parser.parse_args(argv)
    _log.VERBOSE = True
    cfg = yaml.safe_load(open(args.config, "r").read())
    if not torch.cuda.is_available():
        raise RuntimeError("bff_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    device = torch.device("cuda", 0)
    log("Processing class:", args.cls.replace(" ", "_"))
    refine_class(cfg, args.cls, device, ClipTextSimilarity(device))
    return 0
