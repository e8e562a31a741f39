"""TEST INFRASTRUCTURE ONLY -- ctypes front end of oracle/bff_oracle.c (the C restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` leg
may import this. Parity status: PINNED (tests/test_oracle_golden.py, fixtures made by
oracle/make_golden.py from the unmodified reference).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libbff_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "bff_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "_build/libbff_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def set_threads(n: int):
    os.environ["OMP_NUM_THREADS"] = str(n)


def project(xyz, pinv, kmat):
    """-> cam (N,3) f64, pix (N,2) i64. pinv = np.linalg.inv(pose)."""
    xyz, pinv, kmat = _c(xyz, np.float64), _c(pinv, np.float64), _c(kmat, np.float64)
    n = xyz.shape[0]
    cam = np.empty((n, 3), np.float64)
    pix = np.empty((n, 2), np.int64)
    lib().oracle_project(_p(xyz), _p(pinv), _p(kmat), ctypes.c_int64(n), _p(cam), _p(pix))
    return cam, pix


def visibility(cam, pix, depth, thresh):
    cam, pix, depth = _c(cam, np.float64), _c(pix, np.int64), _c(depth, np.float32)
    h, w = depth.shape
    vis = np.empty(cam.shape[0], np.uint8)
    lib().oracle_visibility(_p(cam), _p(pix), _p(depth), ctypes.c_int64(cam.shape[0]), h, w,
                            ctypes.c_double(thresh), _p(vis))
    return vis.astype(bool)


def mask_gather(pix, vis, masks):
    pix, vis, masks = _c(pix, np.int64), _c(vis, np.uint8), _c(masks != 0, np.uint8)
    m, h, w = masks.shape
    rows = np.empty((m, pix.shape[0]), np.uint8)
    lib().oracle_mask_gather(_p(pix), _p(vis), _p(masks), ctypes.c_int64(pix.shape[0]), m, h, w, _p(rows))
    return rows.astype(bool)


def lift_scene(xyz, pinv, kmat, depth, thresh, masks, mask_off, mask_query=None, viewed=None,
               n_query=1, want_rows=True):
    """xyz (N,3); pinv (F,4,4); depth (F,H,W) f32; masks (sumM,H,W) u8; mask_off (F+1).
    -> rows (sumM,N) bool or None, mc (Q,N) i32, vc (N,) i32."""
    xyz, pinv, kmat = _c(xyz, np.float64), _c(pinv, np.float64), _c(kmat, np.float64)
    depth = _c(depth, np.float32)
    f, h, w = depth.shape
    n = xyz.shape[0]
    masks = _c(masks, np.uint8).reshape(-1, h, w)
    mask_off = _c(mask_off, np.int32)
    mq = None if mask_query is None else _c(mask_query, np.int32)
    vw = None if viewed is None else _c(viewed, np.uint8)
    rows = np.empty((masks.shape[0], n), np.uint8) if want_rows else None
    mc = np.empty((n_query, n), np.int32)
    vc = np.empty(n, np.int32)
    lib().oracle_lift_scene(_p(xyz), ctypes.c_int64(n), _p(pinv), _p(kmat), f, _p(depth), h, w,
                            ctypes.c_double(thresh), _p(masks), _p(mask_off), _p(mq), _p(vw),
                            n_query, _p(rows), _p(mc), _p(vc))
    return (rows.astype(bool) if want_rows else None), mc, vc


def mask_iou(a, b):
    """a (nA,N), b (nB,N) bool/u8 -> inter (nA,nB) i32, iou (nA,nB) f32."""
    a, b = _c(a, np.uint8), _c(b, np.uint8)
    na, n = a.shape
    nb = b.shape[0]
    inter = np.empty((na, nb), np.int32)
    iou = np.empty((na, nb), np.float32)
    lib().oracle_mask_iou(_p(a), _p(b), na, nb, ctypes.c_int64(n), _p(inter), _p(iou))
    return inter, iou


def rle_decode(rle):
    counts = _c(np.asarray(rle["counts"]), np.int64)
    length = int(rle["length"])
    m = np.empty(length, np.uint8)
    lib().oracle_rle_decode(_p(counts), ctypes.c_int64(counts.size), ctypes.c_int64(length), _p(m))
    return m


def resize_bilinear(src, height, width):
    """cv2.resize(src, (width, height)) on float32 (projection_2d_to_3d.py:436) restated."""
    src = _c(src, np.float32)
    hs, ws = src.shape
    dst = np.empty((int(height), int(width)), np.float32)
    lib().oracle_resize_bilinear(_p(src), hs, ws, _p(dst), int(height), int(width))
    return dst
