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


def _chain_worker(rank, world, port, q, path):
    """The CSV chain op on two ranks, with the GPU replaced by a stand-in built on the oracle's index:
    what is under test is the slicing of the requested rows and the reduce of the histogram."""
    sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import types

    import torch.distributed as dist
    import csv_oracle
    from ops import _cuda_runtime as rt, map_classify_tpu as op

    C = 4
    data = open(path, "rb").read()

    class FakeIndex(csv_oracle.OracleIndex):
        def classify(self, column, start_row, n_rows, max_len=128, topk=5, want_logits=False):
            blob, fend, _ = self.rows_read(start_row, n_rows, column)
            lens = np.diff(np.concatenate([[0], fend]))
            labels = (lens % C).astype(np.int32)
            top_idx = np.tile(np.arange(topk, dtype=np.int32), (n_rows, 1))
            return {"top_idx": top_idx, "top_score": np.full((n_rows, topk), 0.25, np.float32), "labels": labels,
                    "histogram": np.bincount(labels, minlength=C).astype(np.int64), "logits": None,
                    "metrics": {"items": n_rows, "tokens": int(lens.sum()), "padded_tokens": 128 * n_rows,
                                "batches": 1, "kernel_launches": 92, "h2d_bytes": 0, "d2h_bytes": 0,
                                "gpu_ms_total": 1.0 + rank}}

    fake_index = FakeIndex(data)
    fake_index.info.gpu_ms = 0.5
    fake_index.info.kernel_launches = 6
    engine = types.SimpleNamespace(device=0)
    handle = types.SimpleNamespace(interpreter=engine, output_details=[{"shape": (1, C)}], model_path="fake.b2m")
    op.get_cuda_handle = lambda *a, **k: handle
    rt.get_csv_index = lambda path, eng: (fake_index, data)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        res = op.run({"source_uri": path, "start_row": 3, "shard_size": 101, "topk": 2, "allow_fallback": False})
        q.put((rank, res["shard"], res["labels"], res["label_histogram"], res["row_count"], res["end_row"],
               res["metrics"]["items"], res["metrics"]["world_size"]))
    finally:
        dist.destroy_process_group()


def test_two_rank_csv_chain(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from csv_cases import _table
    path = os.path.join(str(tmp_path), "chain.csv")
    with open(path, "wb") as f:
        f.write(_table(150, 9))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_chain_worker, args=(r, 2, port, q, path)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, s0, l0, h0, n0, e0, i0, w0), (_, s1, l1, h1, n1, e1, i1, w1) = got
    assert (s0["start"], s0["end"], s1["start"], s1["end"]) == (0, 50, 50, 101)
    assert n0 == n1 == 101 and e0 == e1 == 104 and i0 == i1 == 101 and w0 == w1 == 2
    assert h0 == h1 == np.bincount(np.array(l0 + l1), minlength=4).tolist()
    # the two slices together are the text column of rows [3, 104) in order
    import csv as pycsv, io
    rows = list(pycsv.DictReader(io.StringIO(open(path, "rb").read().decode(), newline="")))[3:104]
    assert l0 + l1 == [len(r["text"].encode()) % 4 for r in rows]


def _failing_worker(rank, world, port, q):
    """Rank 1's slice fails; rank 0's is fine. Both must come out of the op with an exception within
    seconds (the failing rank still joins the all-reduce), not one of them stuck in the collective."""
    sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
    import types

    import torch.distributed as dist
    from ops import _shard, map_classify_tpu as op

    C = 4

    def classify(texts, max_len=128, topk=5, want_logits=False):
        if rank == 1:
            raise RuntimeError("device fault on rank 1")
        n = len(texts)
        return {"top_idx": np.zeros((n, topk), np.int32), "top_score": np.zeros((n, topk), np.float32),
                "labels": np.zeros(n, np.int32), "histogram": np.bincount(np.zeros(n, np.int64), minlength=C),
                "logits": None, "metrics": {"items": n, "gpu_ms_total": 1.0}}

    engine = types.SimpleNamespace(device=0, classify=classify)
    handle = types.SimpleNamespace(interpreter=engine, output_details=[{"shape": (1, C)}], model_path="fake.b2m")
    op.get_cuda_handle = lambda *a, **k: handle
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        try:
            op.run({"items": ["a", "b", "c", "d"], "allow_fallback": False})
            q.put((rank, "no exception", ""))
        except Exception as e:  # noqa: BLE001
            q.put((rank, type(e).__name__, str(e)))
        # with the default allow_fallback the same failure becomes the error envelope on every rank
        env = op.run({"items": ["a", "b", "c", "d"]})
        q.put((rank, "envelope", env.get("reason", "")))
    finally:
        dist.destroy_process_group()


def test_one_failing_rank_fails_the_job_on_all_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_failing_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(4))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][:2] == (0, "ShardFailed") and "1 of 2 ranks failed" in got[0][2]
    assert got[1] == (0, "envelope", got[0][2])
    assert got[2] == (1, "RuntimeError", "device fault on rank 1")
    assert got[3] == (1, "envelope", "device fault on rank 1")
