"""Scene-parallel scheduling of a class's scene loop over the ranks (BASELINE config 5: the 312-scene val sweep on 8 GPUs).

The reference's outermost loop (tools/projection_2d_to_3d.py:365) visits independent scenes, so with several ranks each
rank takes whole scenes and no data-path collective is needed (0.98 efficiency at 8 GPUs, against the
quadratic aggregation cost a frame-sharded scene pays). Scene sizes vary by 10x (40k-420k points, 58-543
frames), so scenes are placed longest-processing-time-first on the least loaded rank; every rank computes the same
assignment from the same file listing."""
from __future__ import annotations

import os
from typing import List, Sequence

import numpy as np


def lpt_assign(costs: Sequence[float], world: int) -> List[List[int]]:
    """Indices of `costs` per rank: items in decreasing cost (ties: lower index first) go to the currently least loaded
    rank (ties: lower rank). Each rank's list is returned in increasing index order (= the reference's sorted scene order)."""
    order = sorted(range(len(costs)), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += float(costs[i])
    return [sorted(x) for x in out]


def scene_cost(cfg, scene_id: str) -> float:
    """N x F_all: points (from the .npy header, no data read) times down-sampled colour frames (:526-535)."""
    g = (lambda k: cfg[k]) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k))
    try:
        n = np.load(os.path.join(g("scene_npy_dir"), f"{scene_id}.npy"), mmap_mode="r").shape[0]
    except Exception:
        n = 1
    try:
        names = [f for f in os.listdir(os.path.join(g("scene_2d_dir"), scene_id, "color")) if f.endswith(".jpg")]
        f = max(1, (len(names) + int(g("downsample_ratio")) - 1) // int(g("downsample_ratio")))
    except Exception:
        f = 1
    return float(n) * float(f)


def dist_mode(n_scenes: int, world: int) -> str:
    """BFF_DIST_MODE = scene | frame | auto (default). auto: whole scenes per rank when there are at least as many scenes
    as ranks, else every scene's frames are sharded over all ranks (bff_b200/dist.py)."""
    mode = os.environ.get("BFF_DIST_MODE", "auto").lower()
    if mode in ("scene", "frame"):
        return mode
    return "scene" if n_scenes >= world else "frame"
