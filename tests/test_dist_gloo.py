"""CPU, world_size 2, gloo: the frame-sharded path (bff_b200.dist) -- contiguous frame blocks per rank,
int32 vote all-reduce, rank-ordered row all-gather -- gives exactly the single-process result. The
CUDA kernel is replaced by the oracle through dist.lift_sharded's injection points (host logic only)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import oracle as O
from helpers import pack_rows, random_scene, unpack_rows


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _scene_and_queries():
    from bff_b200 import lift as lift_mod
    from bff_b200.rle import rle_encode_rows
    sc = random_scene(31, n=1500, f=7, h=40, w=56, m_per_frame=(0, 1, 2, 3))
    recs = []
    for f in range(sc["f"]):
        m0, m1 = sc["mask_off"][f], sc["mask_off"][f + 1]
        if m1 > m0:
            recs.append({"frame_id": f"{f * 10}.jpg", "segmented_frame_masks": rle_encode_rows(torch.from_numpy(sc["masks"][m0:m1].reshape(m1 - m0, -1))),
                         "confidences": torch.full((m1 - m0,), 0.3), "labels": ["q"] * (m1 - m0)})
    q = lift_mod.QueryMasks.from_file_records("q", recs)
    view_ids = [str(f * 10) for f in range(sc["f"])]
    return sc, q, view_ids


def _oracle_lift_fn(sc):
    import oracle as O
    from bff_b200 import ops

    def lift_fn(xyz_dev, intrinsics, frames, rles, mask_query, offsets, n_queries, height, width, want_viewed):
        masks = np.stack([O.rle_decode(r).reshape(height, width) for r in rles]) if rles else np.zeros((0, height, width), np.uint8)
        rows, mc, vc = O.lift_scene(sc["xyz"], frames.pose_inv, intrinsics, frames.depth, 0.08, masks,
                                    np.asarray(offsets, np.int32), None, frames.viewed, n_query=1)
        rows_p = torch.from_numpy(pack_rows(rows).view(np.int32)) if rows.shape[0] else torch.zeros((0, (sc["n"] + 31) // 32), dtype=torch.int32)
        return ops.LiftOutput(rows_p, torch.from_numpy(mc), torch.from_numpy(vc))
    return lift_fn


def _load_fn(sc):
    from bff_b200 import scene_io

    def load_fn(ids, viewed):
        idx = [int(i) // 10 for i in ids]
        return scene_io.FrameSet(list(ids), sc["pinv"][idx], sc["depth"][idx], False, np.asarray(viewed, np.uint8), None)
    return load_fn


def _worker(rank, world, port, out_dir):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.join(here, "..", "oracle"), os.path.join(here, "..", "beyond-fixed-forms_b200")):
        sys.path.insert(0, p)
    from bff_b200 import dist as dist_mod, lift as lift_mod
    ctx = dist_mod.init_from_env("gloo")
    sc, q, view_ids = _scene_and_queries()
    frame_ids, viewed, per_frame = lift_mod.plan_frames([q], view_ids)
    lifted = dist_mod.lift_sharded(ctx, sc["xyz"], sc["K"], "", "", frame_ids, viewed, per_frame, [q], sc["h"], sc["w"], True,
                                   torch.device("cpu"), lift_fn=_oracle_lift_fn(sc), load_fn=_load_fn(sc))
    # the all-pairs IoU with one row block per rank (dist.sharded_pairs), oracle injected for the kernel
    import oracle as O
    from helpers import unpack_rows

    def iou_fn(block, rows):
        a, b = unpack_rows(block.numpy(), sc["n"]), unpack_rows(rows.numpy(), sc["n"])
        inter, _ = O.mask_iou(a, b)
        return torch.from_numpy(inter), torch.from_numpy(a.sum(1).astype(np.int32))
    inter, pop = dist_mod.sharded_pairs(ctx, lifted.rows[0], iou_fn=iou_fn)
    torch.save({"rows": lifted.rows[0], "mc": lifted.masked_counts, "vc": lifted.viewed_counts, "inter": inter, "pop": pop},
               os.path.join(out_dir, f"r{rank}.pth"))
    ctx.finalize()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_lift_equals_single(tmp_path, world):
    from bff_b200 import dist as dist_mod, lift as lift_mod
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sc, q, view_ids = _scene_and_queries()
    frame_ids, viewed, per_frame = lift_mod.plan_frames([q], view_ids)
    frames = _load_fn(sc)(frame_ids, viewed)
    rles, mq, offsets = lift_mod.mask_layout(per_frame, [q], 0, len(frame_ids))
    single = _oracle_lift_fn(sc)(None, sc["K"], frames, rles, mq, offsets, 1, sc["h"], sc["w"], True)
    assert single.rows.shape[0] > 3 and int(single.viewed_counts.sum()) > 0
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"r{r}.pth"))
        assert torch.equal(got["rows"], single.rows)
        assert torch.equal(got["mc"], single.masked_counts) and torch.equal(got["vc"], single.viewed_counts)
        dense = unpack_rows(single.rows.numpy(), sc["n"])
        inter_o, _ = O.mask_iou(dense, dense)
        assert np.array_equal(got["inter"].numpy(), inter_o) and np.array_equal(got["pop"].numpy(), dense.sum(1))


def test_shard_bounds_cover_everything():
    from bff_b200.dist import shard_bounds
    for n in (0, 1, 7, 300, 2000):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1
