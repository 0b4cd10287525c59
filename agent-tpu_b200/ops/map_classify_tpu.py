"""map_classify_tpu on B200: same registration, payload and result envelope as the reference op
(ops/map_classify_tpu.py:31-90, ops/map_classify_tpu.CONTRACT.md), batch-capable underneath.

Kept from the reference
  * `@register_op("map_classify_tpu")`, handler `run(payload, ctx=None)`, `payload=None` tolerated;
  * v0 keys: `input` (flat list[int], one pre-tokenised row), `model_path`, `topk` (default 5),
    `allow_fallback` (default True); result `{op, model_path, topk:[{index, score}], elapsed_ms}`;
  * k is clamped to [1, number of classes] (ops/map_classify_tpu.py:16);
  * any exception becomes `{op, fallback:"cpu", reason, topk:[], elapsed_ms}` unless
    `allow_fallback` is falsy, in which case it propagates (ops/map_classify_tpu.py:84-90). The
    "cpu" in that envelope is the reference's wording for "the accelerator did not run"; no CPU
    computation happens there or here.

Added: `items: list[str]` (the key map_tokenize already uses,
ops/map_tokenize.py:29) classifies a whole job: tokenised on the GPU, encoded in batches of
`batch` (default 1024) at padded length `max_len` (default 128). The result then also carries
`items_count`, `labels`, `label_histogram`, `metrics` and, with `return_results`, per-item `results`.
Top-level `topk` stays (first item) so v0 consumers keep working. `source_uri` instead of `items`
runs the chained CSV job (see `_run_csv`): a file reference instead of a million strings in the
payload, which the lease protocol's max_payload_bytes (app.py:108) could not carry.
"""
from __future__ import annotations

import base64
import os
import time
from typing import Any, Dict, List

import numpy as np

from . import register_op
from ._cuda_runtime import get_model_path, get_cuda_handle
from . import _shard

OP_NAME = "map_classify_tpu"
PAD_ID = 0  # [PAD] in BERT vocabularies; a v0 `input` row is masked where it equals this id


def _clamp_k(k: Any, n_classes: int) -> int:
    return max(1, min(int(k), int(n_classes)))


def _topk_dicts(idx_row: np.ndarray, score_row: np.ndarray) -> List[Dict[str, Any]]:
    # plain Python numbers: the result goes through requests.post(json=...) (app.py:146)
    return [{"index": int(i), "score": float(s)} for i, s in zip(idx_row, score_row)]


def _error_envelope(payload: Dict[str, Any]) -> Dict[str, Any]:
    return {
        "op": OP_NAME,
        "fallback": "cpu",
        "reason": payload.get("fallback_reason", "CUDA unavailable"),
        "topk": [],
    }


def _packed_array(a: np.ndarray) -> Dict[str, Any]:
    a = np.ascontiguousarray(a)
    return {"dtype": a.dtype.newbyteorder("<").name, "shape": list(a.shape),
            "data": base64.b64encode(a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes()).decode("ascii")}


def _attach_outputs(result: Dict[str, Any], out: Dict[str, Any], payload: Dict[str, Any], rank: int, world: int) -> None:
    """Per-item outputs in the form the payload asks for.

    default        `labels` (list[int]); `results` / `logits` lists with return_results / return_logits —
                   fine for thousands of items, and what round-1 consumers read.
    result_format  "packed": one base64 string per array (little-endian int32 labels [n], int32 top_index
                   and float32 top_score [n, k], float32 logits [n, C] on request) under `packed`, so a
                   million-item result is ~60 MB of text instead of a JSON list of a million numbers and
                   a million dicts. `unpack_result` turns it back into arrays.
    result_uri     a path: the same arrays go into that .npz file and the result only names it — the
                   output-side twin of `source_uri`, for results beyond what the lease protocol's
                   requests.post(json=...) should carry (app.py:108 max_payload_bytes, app.py:146)."""
    arrays = {"labels": out["labels"], "top_index": out["top_idx"], "top_score": out["top_score"]}
    if payload.get("return_logits") and out.get("logits") is not None:
        arrays["logits"] = out["logits"]
    uri = payload.get("result_uri")
    if uri:
        if not isinstance(uri, str):
            raise ValueError("payload.result_uri must be a path string")
        path = uri if world == 1 else f"{uri}.rank{rank}"
        if not path.endswith(".npz"):
            path += ".npz"
        tmp = path + ".tmp.npz"
        np.savez(tmp, histogram=np.asarray(out["histogram"], dtype=np.int64), **arrays)
        os.replace(tmp, path)
        result["result_uri"] = path
        result["result_bytes"] = os.path.getsize(path)
        result["result_arrays"] = {k: {"dtype": str(v.dtype), "shape": list(v.shape)} for k, v in arrays.items()}
        return
    fmt = payload.get("result_format", "json")
    if fmt == "packed":
        result["packed"] = {"encoding": "base64", "byte_order": "little",
                            **{k: _packed_array(v) for k, v in arrays.items()}}
        return
    if fmt != "json":
        raise ValueError('payload.result_format must be "json" or "packed"')
    result["labels"] = out["labels"].tolist()
    if payload.get("return_results"):
        result["results"] = [{"topk": _topk_dicts(i, s)} for i, s in zip(out["top_idx"], out["top_score"])]
    if payload.get("return_logits"):
        result["logits"] = out["logits"].tolist()


def unpack_result(result: Dict[str, Any]) -> Dict[str, np.ndarray]:
    """Arrays of a result produced with result_format "packed" or result_uri (consumer-side helper)."""
    if "result_uri" in result:
        with np.load(result["result_uri"]) as z:
            return {k: z[k] for k in z.files}
    out = {}
    for k, v in result["packed"].items():
        if isinstance(v, dict):
            out[k] = np.frombuffer(base64.b64decode(v["data"]), dtype=np.dtype(v["dtype"]).newbyteorder("<")
                                   ).reshape(v["shape"])
    return out


def _guarded(work):
    """(result, None) or (None, exception): under torchrun a rank whose slice fails must still reach
    the job's collective (ops/_shard.py: reduce_job), so the failure is carried there, not raised here."""
    try:
        return work(), None
    except Exception as e:  # noqa: BLE001 - re-raised by reduce_job on every rank
        return None, e


def _run_csv(payload: Dict[str, Any], h, eng, k: int, max_len: int, t0: float) -> Dict[str, Any]:
    """The chained job csv_shard -> map_tokenize -> map_classify in one task: `source_uri` names a CSV
    file, `text_column` (default "text") the column holding the items, `start_row` / `shard_size`
    the data rows to take as in read_csv_shard (ops/csv_shard.py:11-12; shard_size defaults to all
    remaining rows). The file is indexed on the GPU and the column goes from the index to the
    tokenizer without leaving the device; under torchrun every rank takes a contiguous slice of
    the rows. Reading problems the reference's reader would hit raise here and so end up in this
    op's error envelope."""
    import os

    from ._cuda_runtime import get_csv_index
    from .csv_shard import reading_failure

    source = payload["source_uri"]
    if not isinstance(source, str) or not source:
        raise ValueError("payload.source_uri must be a non-empty string")
    if not os.path.exists(source):
        raise FileNotFoundError(f"file not found: {source}")
    start_row = int(payload.get("start_row", 0))
    if start_row < 0:
        raise ValueError("payload.start_row must be >= 0")
    index, data = get_csv_index(source, eng)
    first = min(start_row, index.rows)
    shard_size = payload.get("shard_size")
    count = index.rows - first if shard_size is None else min(int(shard_size), index.rows - first)
    if count < 0 or (shard_size is not None and int(shard_size) <= 0):
        raise ValueError("payload.shard_size must be > 0")
    problem = reading_failure(index.info, index, data, first + count)
    if problem:
        raise RuntimeError(f"failed reading csv: {problem}")
    names = index.fieldnames() or []
    column = payload.get("text_column", "text")
    if column not in names:
        raise KeyError(f"column {column!r} is not in the CSV header {names}")
    col = len(names) - 1 - names[::-1].index(column)  # DictReader keeps the last of duplicate names

    lo, hi, rank, world = _shard.my_slice(count)
    out, failure = _guarded(lambda: index.classify(col, first + lo, hi - lo, max_len=max_len, topk=k,
                                                   want_logits=bool(payload.get("return_logits"))))
    metrics = dict(out["metrics"]) if out else {}
    metrics["csv"] = {"bytes": index.n_bytes, "rows": index.rows, "index_gpu_ms": float(index.info.gpu_ms),
                      "index_kernel_launches": int(index.info.kernel_launches)}
    hist, metrics = _shard.reduce_job(out["histogram"] if out else np.zeros(int(h.output_details[0]["shape"][-1]), np.int64),
                                      metrics, eng.device, error=failure)
    result = {
        "op": OP_NAME,
        "model_path": h.model_path,
        "topk": _topk_dicts(out["top_idx"][0], out["top_score"][0]) if hi > lo else [],
        "elapsed_ms": 0.0,
        "dataset_id": payload.get("dataset_id", "unknown_dataset"),
        "start_row": start_row,
        "end_row": start_row + count,
        "row_count": count,
        "items_count": count,
        "label_histogram": [int(v) for v in hist],
        "metrics": metrics,
    }
    if world > 1:
        result["shard"] = {"rank": rank, "world_size": world, "start": lo, "end": hi}
    _attach_outputs(result, out, payload, rank, world)
    result["elapsed_ms"] = (time.time() - t0) * 1000.0
    return result


@register_op(OP_NAME)
def run(payload: Dict[str, Any], ctx: Dict[str, Any] = None) -> Dict[str, Any]:
    """
    B200-backed text classification.

    Uses:
      - payload["model_path"] OR env CUDA_MODEL_PATH / TPU_MODEL_PATH OR /models/model_b200.b2m
      - payload["topk"] default 5
      - payload["items"] (list of texts)  OR  payload["source_uri"] (CSV file, with "text_column",
        "start_row", "shard_size")  OR  payload["input"] (flat list[int], one row of token ids)
      - payload["max_len"] default 128, payload["batch"] default 1024, payload["vocab_path"] optional
      - payload["allow_fallback"] default True
    """
    ctx = ctx or {}
    payload = payload or {}

    t0 = time.time()
    allow_fallback = payload.get("allow_fallback", True)

    try:
        model_path = get_model_path(payload.get("model_path"))
        max_len = int(payload.get("max_len", 128))
        batch = int(payload.get("batch", 1024))
        if max_len <= 0 or batch <= 0:
            raise ValueError("payload.max_len and payload.batch must be positive integers")
        h = get_cuda_handle(model_path, payload.get("vocab_path"), max_len=max_len, max_batch=batch)
        eng = h.interpreter
        n_classes = int(h.output_details[0]["shape"][-1])
        k = _clamp_k(payload.get("topk", 5), n_classes)

        items = payload.get("items")
        if items is not None:
            if not isinstance(items, list):
                raise ValueError("payload.items must be a list of strings")
            texts = ["" if x is None else str(x) for x in items]
            lo, hi, rank, world = _shard.my_slice(len(texts))
            out, failure = _guarded(lambda: eng.classify(texts[lo:hi], max_len=max_len, topk=k,
                                                         want_logits=bool(payload.get("return_logits"))))
            hist, metrics = _shard.reduce_job(out["histogram"] if out else np.zeros(n_classes, np.int64),
                                              out["metrics"] if out else {}, eng.device, error=failure)
            first = _topk_dicts(out["top_idx"][0], out["top_score"][0]) if hi > lo else []
            result = {
                "op": OP_NAME,
                "model_path": h.model_path,
                "topk": first,
                "elapsed_ms": 0.0,
                "items_count": len(texts),
                "label_histogram": [int(v) for v in hist],
                "metrics": metrics,
            }
            if world > 1:
                result["shard"] = {"rank": rank, "world_size": world, "start": lo, "end": hi}
            _attach_outputs(result, out, payload, rank, world)
            result["elapsed_ms"] = (time.time() - t0) * 1000.0
            return result

        if payload.get("source_uri") is not None:
            return _run_csv(payload, h, eng, k, max_len, t0)

        if "input" not in payload:
            raise ValueError('payload missing required key: "input"')

        input_details = h.input_details[0]
        shape = (1, max_len)
        dtype = input_details["dtype"]

        arr = np.array(payload["input"], dtype=dtype)

        expected = int(np.prod(shape))
        if arr.size != expected:
            raise ValueError(
                f"Input size mismatch. Got {arr.size}, expected {expected} for shape {shape}."
            )

        arr = arr.reshape(shape)
        out = eng.classify_ids(arr, (arr != PAD_ID).astype(np.uint8), topk=k)

        return {
            "op": OP_NAME,
            "model_path": h.model_path,
            "topk": _topk_dicts(out["top_idx"][0], out["top_score"][0]),
            "elapsed_ms": (time.time() - t0) * 1000.0,
        }

    except Exception as e:
        if allow_fallback:
            return {
                **_error_envelope({"fallback_reason": str(e), **payload}),
                "elapsed_ms": (time.time() - t0) * 1000.0,
            }
        raise
