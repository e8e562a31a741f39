"""Concurrent per-query work on one GPU.

After the shared lifting launch every text query is aggregated, filtered and refined on its own rows
(tools/projection_2d_to_3d.py:463-634 and tools/refinement.py:186-412 are per --cls in the reference). One query
is a chain of small kernels separated by host decisions, so a single chain leaves the GPU idle most of the time;
QueryPool runs several chains at once, each on its own CUDA stream with its own bff_ctx scratch arenas (the C
pipelines release the GIL while they wait for the device). Results are identical to the sequential loop: the
chains share only read-only inputs."""
from __future__ import annotations

import threading
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Iterable, List

import torch


def _cuda_tensors(obj):
    if isinstance(obj, torch.Tensor):
        if obj.is_cuda:
            yield obj
    elif isinstance(obj, dict):
        for v in obj.values():
            yield from _cuda_tensors(v)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            yield from _cuda_tensors(v)


class QueryPool:
    def __init__(self, device, workers: int = 2):
        self.device = torch.device(device)
        self.workers = max(1, int(workers))
        self._tls = threading.local()
        self._exec = ThreadPoolExecutor(max_workers=self.workers, thread_name_prefix="bff-query") if self.workers > 1 else None

    def _run(self, fn: Callable, item, ready: torch.cuda.Event):
        if not hasattr(self._tls, "stream"):
            torch.cuda.set_device(self.device)
            self._tls.stream = torch.cuda.Stream(device=self.device)
        s = self._tls.stream
        with torch.cuda.stream(s):
            s.wait_event(ready)                  # inputs were produced on the caller's stream
            out = fn(item)
            done = torch.cuda.Event()
            done.record(s)
        return out, done

    def map(self, fn: Callable, items: Iterable) -> List:
        """[fn(item) for item in items], the calls spread over the worker streams. On return the caller's current
        stream has been made to wait for every worker, and the CUDA tensors in the results are safe to use on it."""
        items = list(items)
        if self._exec is None or len(items) <= 1:
            return [fn(it) for it in items]
        main = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(main)
        futs = [self._exec.submit(self._run, fn, it, ready) for it in items]
        outs = []
        for fu in futs:
            out, done = fu.result()
            main.wait_event(done)
            for t in _cuda_tensors(out):
                t.record_stream(main)            # allocated on a worker stream, consumed on the caller's
            outs.append(out)
        return outs

    def close(self):
        if self._exec is not None:
            self._exec.shutdown(wait=True)
            self._exec = None


_pools = {}


def get_pool(device, workers: int = 2) -> QueryPool:
    key = (str(torch.device(device)), int(workers))
    if key not in _pools:
        _pools[key] = QueryPool(device, workers)
    return _pools[key]

# Measured on a B200 (bench.py --workload c3: 20 queries, 1351 lifted masks): 17.9 ms per scene with the sequential loop,
# 14.3 ms with 2 workers, 19.0 with 4, 17.1 with 8; with two ranks on one host (2 workers each) 27 ms. Beyond the C calls
# the Python glue of the chains contends for the GIL, so the gain is small and not stable: callers default to 1 worker
# (BFF_QUERY_WORKERS / stage2.QUERY_WORKERS opt in). The default for several queries is the batched pipelines instead
# (bff_stage2_aggregate_batch / bff_refine_match_batch: one C++ call drives every query's phases, 6.9 ms per scene at c3).
