"""The N>1 path on CPU: two processes, gloo, contiguous slices and the per-job histogram/metrics
all-reduce (ops/_shard.py) — the same code NCCL runs on the GPUs."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
    import torch.distributed as dist
    from ops import _shard
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        n = 1001
        lo, hi, r, w = _shard.my_slice(n)
        labels = np.arange(lo, hi) % 5                       # stand-in for this rank's arg-max labels
        hist = np.bincount(labels, minlength=5).astype(np.int64)
        metrics = {"items": hi - lo, "tokens": 10 * (hi - lo), "padded_tokens": 128 * (hi - lo), "batches": 1,
                   "kernel_launches": 90, "h2d_bytes": 1, "d2h_bytes": 2, "gpu_ms_total": 5.0 + rank}
        total, merged = _shard.reduce_job(hist, metrics)
        q.put((rank, lo, hi, total.tolist(), merged))
    finally:
        dist.destroy_process_group()


def test_two_rank_histogram_allreduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, h0, m0), (r1, lo1, hi1, h1, m1) = got
    assert (lo0, hi0, lo1, hi1) == (0, 500, 500, 1001)
    want = np.bincount(np.arange(1001) % 5, minlength=5).tolist()
    assert h0 == want and h1 == want
    for m in (m0, m1):
        assert m["items"] == 1001 and m["tokens"] == 10010 and m["kernel_launches"] == 180
        assert m["gpu_ms_total"] == 6.0 and m["world_size"] == 2
