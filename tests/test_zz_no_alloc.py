"""GPU: bff_vote_unique_static (include/bff.h) -- the allocation-free variant of bff_vote_unique whose scratch hash table
is a file-scope device array inside libbff.so -- called through the C ABI and compared with the existing entry point.

Collected last on purpose (a failure here must not hide another test). Bit-exact like the
other vote tests: both entry points run the same kernels, only the scratch table's home differs. Two calls in a row share
the table, so the second (smaller, different values) would show stale slots if the call did not re-initialise them."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _hist(ops, dev, seed, n, mc_hi, vc_hi, mode):
    rng = np.random.default_rng(seed)
    mc = rng.integers(0, mc_hi, n).astype(np.int32)
    mc[rng.random(n) < 0.4] = 0
    vc = rng.integers(0, vc_hi, n).astype(np.int32)
    mc_t, vc_t = torch.from_numpy(mc).to(dev), torch.from_numpy(vc).to(dev)
    hist = ops.vote_hist(mc_t, vc_t if mode == 0 else None, int(mc.max()) + 1, int(vc.max()) + 1 if mode == 0 else 1)
    mcf, vcf = torch.from_numpy(mc).float(), torch.from_numpy(vc).float()
    return hist, ((mcf / (vcf + 1)).unique() if mode == 0 else mcf.unique())


def _static_call(_cabi, hist, mode, values, count):
    rc = _cabi.lib().bff_vote_unique_static(_cabi.ptr(hist), hist.shape[0], hist.shape[1], mode, _cabi.ptr(values), _cabi.ptr(count),
                                            _cabi.current_stream())
    _cabi.check(rc, "bff_vote_unique_static")
    return torch.sort(values[: int(count.item())]).values


def test_static_scratch_entry_point_equals_the_existing_one_on_two_calls_in_a_row(cuda_dev):
    from bff_b200 import _cabi, ops
    # caller-owned outputs, allocated once up front: the calls below must not need anything else
    values = torch.empty((_cabi.VOTE_UNIQUE_STATIC_MAX_BINS,), dtype=torch.float32, device=cuda_dev)
    count = torch.zeros((1,), dtype=torch.int32, device=cuda_dev)
    cases = [(0, 31, 200_003, 140, 301),      # 42 140 bins: a 128 Ki-word table
             (0, 32, 50_001, 9, 40),          # right after it: 360 bins, the smallest table -- stale slots would add values
             (1, 33, 80_000, 500, 7),         # masked_counts.unique()
             (0, 34, 120_000, 400, 301)]      # 120 400 bins: the largest table the static array holds (256 Ki words)
    for mode, seed, n, mc_hi, vc_hi in cases:
        hist, want = _hist(ops, cuda_dev, seed, n, mc_hi, vc_hi, mode)
        assert hist.numel() <= _cabi.VOTE_UNIQUE_STATIC_MAX_BINS
        existing = ops.vote_unique(hist, mode)                       # caller-provided table (the existing entry point)
        got = _static_call(_cabi, hist, mode, values, count)
        assert torch.equal(got, existing), (mode, seed)
        assert torch.equal(got.cpu(), want) and len(want) > 5        # and both are the reference's unique()
    # the same input twice in a row: identical, and nothing was allocated by torch or the driver in between
    hist, want = _hist(ops, cuda_dev, 35, 60_000, 60, 200, 0)
    first = _static_call(_cabi, hist, 0, values, count)
    torch.cuda.synchronize()
    torch_before, free_before = torch.cuda.memory_allocated(cuda_dev), torch.cuda.mem_get_info(cuda_dev)[0]
    ops.vote_unique_static(hist, 0, values, count)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(cuda_dev) == torch_before
    assert free_before - torch.cuda.mem_get_info(cuda_dev)[0] < (1 << 20)          # no cudaMalloc (the table alone is 1 MiB)
    second = torch.sort(values[: int(count.item())]).values
    assert torch.equal(first, second) and torch.equal(second.cpu(), want)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_table(cuda_dev):
    from bff_b200 import _cabi
    mb, vb = 700, 301                                               # 210 700 bins > 131 072
    assert mb * vb > _cabi.VOTE_UNIQUE_STATIC_MAX_BINS
    hist = torch.zeros((mb, vb), dtype=torch.int32, device=cuda_dev)
    values = torch.empty((mb * vb,), dtype=torch.float32, device=cuda_dev)
    count = torch.full((1,), -7, dtype=torch.int32, device=cuda_dev)
    rc = _cabi.lib().bff_vote_unique_static(_cabi.ptr(hist), mb, vb, 0, _cabi.ptr(values), _cabi.ptr(count), _cabi.current_stream())
    assert rc != 0
    with pytest.raises(_cabi.BffError, match="limit of 131072 bins"):
        _cabi.check(rc, "bff_vote_unique_static")
    torch.cuda.synchronize()
    assert int(count.item()) == -7                                   # rejected before anything was enqueued
    # exactly at the limit is accepted (an empty histogram has no values)
    hist = torch.zeros((512, 256), dtype=torch.int32, device=cuda_dev)
    rc = _cabi.lib().bff_vote_unique_static(_cabi.ptr(hist), 512, 256, 1, _cabi.ptr(values), _cabi.ptr(count), _cabi.current_stream())
    _cabi.check(rc, "bff_vote_unique_static")
    assert int(count.item()) == 0
