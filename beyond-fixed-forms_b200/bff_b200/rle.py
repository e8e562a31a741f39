"""Run-length wire format of 2D masks and stage-1 masks (host side).

Format (tools/utils/rle_encode_decode.py:10-33): ``{"length": L, "counts": int array of
(1-based start, run length) pairs}`` over the flattened mask. Decoding semantics follow
rle_encode_decode.py:35-61 / tools/refinement.py:26-39: pairs are cast to int32, start is
made 0-based, and each run is written with Python slice semantics (``mask[lo:hi] = 1``),
so runs may overlap and out-of-range ends are clipped.

The product path never expands a mask to one byte per element: `runs_for_device` turns the
pairs into clipped [lo, hi) bit ranges and the `bff_rle_to_bits` kernel sets the bits of
the packed row directly.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np
import torch


def rle_encode_rows(masks: torch.Tensor) -> List[dict]:
    """Encode (n, L) boolean rows. Same output as rle_encode_batch (rle_encode_decode.py:10-33):
    positions where the zero-padded row changes value, odd entries turned into lengths."""
    m = torch.as_tensor(masks).reshape(masks.shape[0], -1).to(torch.bool).cpu().numpy()
    n, L = m.shape
    out = []
    for i in range(n):
        padded = np.zeros(L + 2, np.int8)
        padded[1:-1] = m[i]
        edges = np.flatnonzero(padded[1:] != padded[:-1]) + 1
        edges[1::2] -= edges[0::2]
        out.append({"length": L, "counts": edges.astype(np.int64)})
    return out


def _slice_bounds(lo: np.ndarray, hi: np.ndarray, L: int) -> Tuple[np.ndarray, np.ndarray]:
    """Python ``a[lo:hi]`` index normalisation for arrays of bounds."""
    lo = np.where(lo < 0, np.maximum(lo + L, 0), np.minimum(lo, L))
    hi = np.where(hi < 0, np.maximum(hi + L, 0), np.minimum(hi, L))
    return lo, hi


def runs_of(rle: dict) -> Tuple[np.ndarray, np.ndarray, int]:
    """Clipped 0-based [lo, hi) ranges of one RLE dict (int64 arrays) and its length."""
    L = int(rle["length"])
    s = np.asarray(rle["counts"])
    if isinstance(rle["counts"], torch.Tensor):
        s = rle["counts"].cpu().numpy()
    starts = np.asarray(s[0::2], dtype=np.int32).astype(np.int64) - 1
    nums = np.asarray(s[1::2], dtype=np.int32).astype(np.int64)
    k = min(starts.size, nums.size)          # zip() drops an unpaired trailing start
    starts, nums = starts[:k], nums[:k]
    ends = (starts + nums).astype(np.int32).astype(np.int64)  # int32 arithmetic in the reference
    lo, hi = _slice_bounds(starts, ends, L)
    keep = hi > lo
    return lo[keep], hi[keep], L


def runs_for_device(rles: Sequence[dict], expect_length: int | None = None):
    """Concatenate the runs of many RLE dicts for `bff_rle_to_bits`.

    Returns (runs int32 (R,3) = [row, lo, hi), length). All rows must share one length."""
    rows, los, his = [], [], []
    L = expect_length
    for r, rle in enumerate(rles):
        lo, hi, Lr = runs_of(rle)
        if L is None:
            L = Lr
        if Lr != L:
            raise ValueError(f"RLE row {r} has length {Lr}, expected {L}")
        rows.append(np.full(lo.shape, r, np.int64))
        los.append(lo)
        his.append(hi)
    if L is None:
        L = 0
    if rows:
        runs = np.stack([np.concatenate(rows), np.concatenate(los), np.concatenate(his)], axis=1)
    else:
        runs = np.zeros((0, 3), np.int64)
    if L >= 2 ** 31:
        raise ValueError("mask length must fit int32")
    return np.ascontiguousarray(runs.astype(np.int32)), int(L)


def pairs_for_device(rles: Sequence[dict], expect_length: int | None = None):
    """The wire format as `bff_rle_pairs_to_bits` takes it: (counts int64 (2P,), pair_offset int64 (R+1,), length). Only
    array concatenation happens here -- the int32 casts and the slice clipping of the reference's decoder run in the
    kernel -- so the host cost per scene is one np.concatenate instead of one numpy pipeline per mask."""
    L = expect_length
    parts, n_pairs = [], np.zeros(len(rles) + 1, np.int64)
    for r, rle in enumerate(rles):
        Lr = int(rle["length"])
        if L is None:
            L = Lr
        if Lr != L:
            raise ValueError(f"RLE row {r} has length {Lr}, expected {L}")
        c = rle["counts"]
        c = c.cpu().numpy() if isinstance(c, torch.Tensor) else np.asarray(c)
        k = c.shape[0] // 2                      # zip() drops an unpaired trailing start
        parts.append(c[:2 * k])
        n_pairs[r + 1] = k
    if L is None:
        L = 0
    if L >= 2 ** 31:
        raise ValueError("mask length must fit int32")
    counts = np.concatenate(parts).astype(np.int64, copy=False) if parts else np.zeros(0, np.int64)
    return np.ascontiguousarray(counts), np.cumsum(n_pairs), int(L)
