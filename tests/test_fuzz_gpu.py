"""GPU: randomized parity sweep of bff_lift_scene against the C oracle -- ragged point counts, odd image sizes, every
cloud / depth representation, several queries, random chunking, culling windows on and off, depth at mask size and at a
different native size (in-kernel resize). Seeds are fixed: failures reproduce."""
import numpy as np
import pytest
import torch

from helpers import pack_rows, random_scene, unpack_rows

pytestmark = pytest.mark.gpu


def _case(seed):
    rng = np.random.default_rng(10_000 + seed)
    n = int(rng.choice([1, 31, 32, 33, 257, 1000, 4099, 9001]))
    f = int(rng.integers(1, 41))
    h, w = int(rng.integers(17, 97)), int(rng.integers(23, 131))
    nq = int(rng.choice([1, 1, 2, 5]))
    return dict(n=n, f=f, h=h, w=w, nq=nq, f32=bool(rng.integers(0, 2)), kind=str(rng.choice(["f32", "u16", "bilinear"])),
                fpc=int(rng.choice([0, 0, 1, 5, 32, 64])), window=bool(rng.integers(0, 2)), viewed=bool(rng.integers(0, 2)))


@pytest.mark.parametrize("seed", range(36))
def test_random_lift_configuration_matches_oracle(cuda_dev, seed):
    import oracle as O
    from bff_b200 import ops
    c = _case(seed)
    sc = random_scene(500 + seed, n=c["n"], f=c["f"], h=c["h"], w=c["w"], n_queries=c["nq"], f32_xyz=c["f32"])
    rng = np.random.default_rng(seed)
    native = None
    if c["kind"] == "bilinear":
        hs, ws = int(rng.integers(9, 80)), int(rng.integers(11, 100))
        raw = np.stack([np.clip(np.round(O.resize_bilinear(d, hs, ws).astype(np.float64) * 1000), 0, 65535) for d in sc["depth"]]).astype(np.uint16)
        sc["depth"] = np.stack([O.resize_bilinear(r.astype(np.float32) / 1000, c["h"], c["w"]) for r in raw])
        depth_dev, native = torch.from_numpy(raw.view(np.int16)).to(cuda_dev), (hs, ws)
    elif c["kind"] == "u16":
        raw = np.round(sc["depth"].astype(np.float64) * 1000).astype(np.uint16)
        sc["depth"] = raw.astype(np.float32) / 1000
        depth_dev = torch.from_numpy(raw.view(np.int16)).to(cuda_dev)
    else:
        depth_dev = torch.from_numpy(sc["depth"]).to(cuda_dev)
    viewed = (rng.random(c["f"]) < 0.6).astype(np.uint8) if c["viewed"] else None
    rows_o, mc_o, vc_o = O.lift_scene(sc["xyz"], sc["pinv"], sc["K"], sc["depth"], 0.08, sc["masks"].astype(np.uint8), sc["mask_off"],
                                      sc["mask_query"] if c["nq"] > 1 else None, viewed, n_query=c["nq"])
    if c["f32"]:
        x = np.zeros((c["n"], 4), np.float32)
        x[:, :3] = sc["xyz"].astype(np.float32)
    else:
        x = sc["xyz"]
    nz = sc["depth"][sc["depth"] > 0]
    window = (0.0 if c["kind"] == "bilinear" else float(nz.min()), float(nz.max())) if (c["window"] and nz.size) else None
    n_masks = sc["masks"].shape[0]
    packed2d = pack_rows(sc["masks"].reshape(n_masks, -1)).view(np.int32) if n_masks else None
    out = ops.lift_scene(torch.from_numpy(x).to(cuda_dev), torch.from_numpy(sc["pinv"]).to(cuda_dev), sc["K"], depth_dev, c["h"], c["w"],
                         masks2d=torch.from_numpy(packed2d).to(cuda_dev) if n_masks else None,
                         mask_offset=torch.from_numpy(sc["mask_off"]).to(cuda_dev) if n_masks else None,
                         mask_query=torch.from_numpy(sc["mask_query"]).to(cuda_dev) if (n_masks and c["nq"] > 1) else None,
                         n_queries=c["nq"], frame_viewed=None if viewed is None else torch.from_numpy(viewed).to(cuda_dev),
                         depth_thresh=0.08, depth_range=window, frames_per_chunk=c["fpc"], native_size=native)
    np.testing.assert_array_equal(unpack_rows(out.rows.cpu().numpy(), c["n"]) if n_masks else np.zeros((0, c["n"]), bool), rows_o, err_msg=str(c))
    np.testing.assert_array_equal(out.masked_counts.cpu().numpy(), mc_o, err_msg=str(c))
    np.testing.assert_array_equal(out.viewed_counts.cpu().numpy(), vc_o, err_msg=str(c))
