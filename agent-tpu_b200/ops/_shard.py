"""Contiguous-slice sharding of a job's items across the GPUs of one box.

One process per GPU (torchrun): rank r of W takes items [r*N//W, (r+1)*N//W). The data path has no
collective — items are independent; the one exchange per job is a sum all-reduce (NCCL over
NVLink) of the label histogram and the additive metrics. Outside torch.distributed this module is
inert: one slice, no reduction, and torch is never imported.
"""
from __future__ import annotations

import sys
from typing import Dict, Optional, Tuple

import numpy as np

_ADDITIVE = ("items", "tokens", "padded_tokens", "batches", "kernel_launches", "h2d_bytes", "d2h_bytes")


def _dist():
    td = sys.modules.get("torch.distributed")
    if td is not None and td.is_available() and td.is_initialized() and td.get_world_size() > 1:
        return td
    return None


def slice_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    return (rank * n) // world, ((rank + 1) * n) // world


def my_slice(n: int) -> Tuple[int, int, int, int]:
    td = _dist()
    if td is None:
        return 0, n, 0, 1
    rank, world = td.get_rank(), td.get_world_size()
    lo, hi = slice_bounds(n, rank, world)
    return lo, hi, rank, world


class ShardFailed(RuntimeError):
    """Another rank's slice of the job failed; this rank's own slice was fine."""


def reduce_job(hist: np.ndarray, metrics: Dict, device: int = 0, error: Optional[BaseException] = None):
    """Sum the label histogram and additive counters over ranks; gpu_ms_total becomes the max.

    Every rank must call this exactly once per job, also when its own slice failed (pass the
    exception as `error`, with a zero histogram): the vector carries a failure count, so that a
    rank that failed still enters the collective instead of leaving the others waiting in it, and
    all ranks then raise together — the failing ones their own exception, the rest ShardFailed."""
    td = _dist()
    if td is None:
        if error is not None:
            raise error
        return hist, metrics
    import torch

    backend = td.get_backend()
    dev = torch.device("cuda", device) if backend == "nccl" else torch.device("cpu")
    vec = np.concatenate([np.asarray(hist, dtype=np.int64),
                          np.array([int(metrics.get(k, 0)) for k in _ADDITIVE], dtype=np.int64),
                          np.array([0 if error is None else 1], dtype=np.int64)])
    t = torch.from_numpy(vec).to(dev)
    td.all_reduce(t, op=td.ReduceOp.SUM)
    failures = int(t[-1].item())
    if failures:
        # the second collective below is skipped by every rank alike
        if error is not None:
            raise error
        raise ShardFailed(f"{failures} of {td.get_world_size()} ranks failed on their slice of the job")
    ms = torch.tensor([float(metrics.get("gpu_ms_total", 0.0))], dtype=torch.float64, device=dev)
    td.all_reduce(ms, op=td.ReduceOp.MAX)
    out = t.cpu().numpy()
    merged = dict(metrics)
    for i, k in enumerate(_ADDITIVE):
        merged[k] = int(out[len(hist) + i])
    merged["gpu_ms_total"] = float(ms.item())
    merged["world_size"] = td.get_world_size()
    return out[: len(hist)], merged
