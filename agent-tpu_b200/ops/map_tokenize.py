"""map_tokenize on B200 (reference ops/map_tokenize.py:13-61).

Default mode reproduces the reference exactly: text is cut into chunks of `chunk_size` code points
(default 1024); `items` are flattened, `None` -> "", anything else -> `str(x)`; the result is
`{ok, tokens, count, total_chars[, items_count]}`; bad input returns `{"ok": False, "error": ...}`
and never raises. Code points are counted and the chunk boundaries found on the GPU
(b2_chunk_count / b2_chunk_bounds); Python only slices the UTF-8 buffer at those byte offsets.

`mode: "wordpiece"` (an addition — the reference has no vocabulary) runs the same
WordPiece tokenize-and-pad kernel the classifier uses and returns `input_ids` / `attention_mask`
rows of length `max_len`.
"""
from typing import Any, Dict, List

import numpy as np

from . import register_op
from ._cuda_runtime import get_tokenizer_engine, pack_items


def _gpu_chunks(strings: List[str], chunk_size: int):
    """(chunks, total code points) for `strings`, boundaries computed on the GPU."""
    blob, offsets = pack_items(strings)
    if blob.size == 0:
        return [], 0
    eng = get_tokenizer_engine(None, max_len=2)
    counts = eng.chunk_counts(blob, offsets)
    per_item = -(-counts // chunk_size)
    base = np.zeros(len(strings), dtype=np.int64)
    np.cumsum(per_item[:-1], out=base[1:])
    n_chunks = int(per_item.sum())
    starts = eng.chunk_bounds(blob, offsets, chunk_size, base, n_chunks)
    ends = np.empty(n_chunks, dtype=np.int64)
    ends[:-1] = starts[1:]
    has = per_item > 0
    ends[(base + per_item - 1)[has]] = offsets[1:][has]
    raw = blob.tobytes()
    chunks = [raw[b:e].decode("utf-8", "surrogatepass") for b, e in zip(starts.tolist(), ends.tolist())]
    return chunks, int(counts.sum())


@register_op("map_tokenize")
def map_tokenize_op(payload: Dict[str, Any]) -> Dict[str, Any]:
    """
    Slices text into fixed-size chunks (default 1024 code points).
    Supports either:
      - {"text": "..."} or {"data": "..."}
      - {"items": ["...", "..."], "chunk_size": 1024}
      - {"items": [...], "mode": "wordpiece", "max_len": 128, "vocab_path": "..."}
    """
    if payload is None:
        payload = {}

    chunk_size = payload.get("chunk_size", 1024)
    if not isinstance(chunk_size, int) or chunk_size <= 0:
        return {"ok": False, "error": "payload.chunk_size must be a positive integer"}

    mode = payload.get("mode", "chunk")
    if mode not in ("chunk", "wordpiece"):
        return {"ok": False, "error": "payload.mode must be 'chunk' or 'wordpiece'"}

    single = None
    if "items" in payload and payload["items"] is not None:
        items = payload["items"]
        if not isinstance(items, list):
            return {"ok": False, "error": "payload.items must be a list of strings"}
        strings = ["" if x is None else str(x) for x in items]
    else:
        single = payload.get("text") or payload.get("data", "")
        if not isinstance(single, str):
            return {"ok": False, "error": "payload.text must be a string"}
        strings = [single]

    try:
        if mode == "wordpiece":
            max_len = payload.get("max_len", 128)
            if not isinstance(max_len, int) or max_len < 2:
                return {"ok": False, "error": "payload.max_len must be an integer >= 2"}
            vocab_path = payload.get("vocab_path")
            if not vocab_path:
                import os
                vocab_path = os.environ.get("CUDA_VOCAB_PATH")
            if not vocab_path:
                return {"ok": False, "error": "payload.vocab_path (or $CUDA_VOCAB_PATH) is required for mode 'wordpiece'"}
            eng = get_tokenizer_engine(vocab_path, max_len=max_len)
            ids, mask, _ = eng.tokenize(strings, max_len=max_len)
            return {
                "ok": True,
                "mode": "wordpiece",
                "input_ids": ids.tolist(),
                "attention_mask": mask.tolist(),
                "count": int(mask.sum()),
                "items_count": len(strings),
                "max_len": max_len,
            }
        chunks, total_chars = _gpu_chunks(strings, chunk_size)
    except Exception as e:  # no CUDA device / library not built: report, never raise (reference style)
        return {"ok": False, "error": f"map_tokenize: {type(e).__name__}: {e}"}

    out = {
        "ok": True,
        "tokens": chunks,
        "count": len(chunks),
        "total_chars": total_chars,
    }
    if single is None:
        out["items_count"] = len(strings)
    return out
