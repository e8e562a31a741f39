"""Shared builders of seeded test inputs (small scenes in memory)."""
import numpy as np


def look_at(eye, target):
    fwd = target - eye
    fwd = fwd / np.linalg.norm(fwd)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return c2w


def pack_rows(dense):
    """(R,N) bool -> (R, ceil(N/32)) uint32, bit b of word w = element 32*w+b (numpy reference)."""
    dense = np.asarray(dense, dtype=bool)
    r, n = dense.shape
    wd = (n + 31) // 32
    pad = np.zeros((r, wd * 32), dtype=bool)
    pad[:, :n] = dense
    bits = np.packbits(pad.reshape(r, wd, 32), axis=2, bitorder="little")
    return bits.reshape(r, wd, 4).copy().view("<u4").reshape(r, wd)


def unpack_rows(packed, n):
    packed = np.ascontiguousarray(packed).view("<u4")
    r = packed.shape[0]
    bits = np.unpackbits(packed.view(np.uint8).reshape(r, -1), axis=1, bitorder="little")
    return bits[:, :n].astype(bool)


def random_scene(seed, n=5000, f=7, h=60, w=80, m_per_frame=(0, 1, 2, 3), n_queries=1, f32_xyz=True):
    """A box room seen by a few cameras with a depth image that matches a good part of the
    points (so that visibility is non-trivial), random blob masks, plus hostile values."""
    rng = np.random.default_rng(seed)
    xyz = rng.uniform([0, 0, 0], [6, 5, 2.8], size=(n, 3))
    if f32_xyz:
        xyz = xyz.astype(np.float32).astype(np.float64)
    fx = 577.870605 * w / 640.0
    K = np.array([[fx, 0.0, (w - 1) / 2.0], [0.0, fx, (h - 1) / 2.0], [0.0, 0.0, 1.0]])
    poses = np.empty((f, 4, 4))
    depth = np.zeros((f, h, w), np.float32)
    for i in range(f):
        eye = rng.uniform([1, 1, 1], [5, 4, 1.8])
        tgt = rng.uniform([0, 0, 0], [6, 5, 2])
        poses[i] = look_at(eye, tgt)
        pinv = np.linalg.inv(poses[i])
        cam = (pinv @ np.concatenate([xyz, np.ones((n, 1))], 1).T).T[:, :3]
        with np.errstate(all="ignore"):
            uv = (K @ cam.T / cam[:, 2]).T[:, :2]
        px = np.round(uv).astype(np.int64)
        ok = (cam[:, 2] > 0.1) & (px[:, 0] >= 0) & (px[:, 0] < w) & (px[:, 1] >= 0) & (px[:, 1] < h)
        # z-buffer of a random half of the points, quantised to millimetres like a sensor
        sel = np.nonzero(ok & (rng.random(n) < 0.5))[0]
        order = sel[np.argsort(-cam[sel, 2])]
        dimg = np.zeros((h, w))
        dimg[px[order, 1], px[order, 0]] = cam[order, 2]
        dimg = np.round(dimg * 1000.0)
        dimg += rng.integers(-90, 91, size=dimg.shape) * (dimg > 0)     # straddle the 0.08 m threshold
        dimg[rng.random((h, w)) < 0.05] = 0
        depth[i] = (np.clip(dimg, 0, 65535).astype(np.uint16).astype(np.float32) / 1000)
    masks, off, query = [], [0], []
    yy, xx = np.mgrid[0:h, 0:w]
    for i in range(f):
        m = int(rng.choice(m_per_frame))
        for _ in range(m):
            cy, cx, r = rng.uniform(0, h), rng.uniform(0, w), rng.uniform(3, max(h, w) / 2.5)
            masks.append(((yy - cy) ** 2 + (xx - cx) ** 2) < r * r)
            query.append(int(rng.integers(0, n_queries)))
        off.append(len(masks))
    masks = np.stack(masks) if masks else np.zeros((0, h, w), bool)
    return dict(xyz=xyz, K=K, poses=poses, pinv=np.stack([np.linalg.inv(p) for p in poses]), depth=depth,
                masks=masks, mask_off=np.asarray(off, np.int32), mask_query=np.asarray(query, np.int32),
                h=h, w=w, n=n, f=f, n_queries=n_queries)
