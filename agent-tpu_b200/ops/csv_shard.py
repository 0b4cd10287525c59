"""`read_csv_shard` on the B200: a slice of a CSV file's data rows, cut from a record index that a
CUDA scan builds once per file (csrc/b2_csv.cu), instead of re-parsing the file from its first
byte with `csv.DictReader` for every shard (reference ops/csv_shard.py:9-26).

Same envelope as the reference op (ops/csv_shard.py:29-103):
  payload   dataset_id (optional), source_uri (required), start_row (default 0, 0-based over data
            rows), shard_size (default 100), mode "rows" | "count"; either the payload itself or a
            task dict carrying it under "payload";
  result    {"ok": True, dataset_id, mode, start_row, end_row, row_count[, rows]} or
            {"ok": False, "error": "read_csv_shard: ..."} — this op never raises.

The reference registers the handler as "read_csv_shard" while its name table lists the module under
"csv_shard", so neither name resolves there; here both names reach this handler.

There is no CPU parser behind this op: without the CUDA library and a device it answers with the
error envelope.
"""
from __future__ import annotations

import codecs
import os
from typing import Any, Dict, List, Optional

from . import register_op

OP = "read_csv_shard"
_READ_CHUNK = 8192  # io.TextIOWrapper decodes the file in reads of this many bytes


def _fail(text: str) -> Dict[str, Any]:
    return {"ok": False, "error": f"{OP}: {text}"}


def _line_end_after(data, pos: int) -> int:
    """Offset of the last byte of the physical line that holds byte `pos` (the reader pulls whole
    lines: "\\n", "\\r" and "\\r\\n" all end one)."""
    n = len(data)
    a, b = data.find(b"\n", pos), data.find(b"\r", pos)
    hits = [x for x in (a, b) if x >= 0]
    if not hits:
        return n - 1
    e = min(hits)
    if data[e:e + 1] == b"\r" and e + 1 < n:
        e += 1  # the reader has to look at the next byte to tell "\r" from "\r\n"
    return e


def _decode_error_text(data, first_bad: int) -> str:
    """The reference's text for a strict UTF-8 failure: replay the 8 KiB read that trips over it
    through Python's own incremental decoder, primed with the read before it (which may end inside
    a sequence)."""
    k = min(first_bad, max(len(data) - 1, 0)) // _READ_CHUNK
    dec = codecs.getincrementaldecoder("utf-8")()
    try:
        if k > 0:
            dec.decode(bytes(data[(k - 1) * _READ_CHUNK:k * _READ_CHUNK]))
        dec.decode(bytes(data[k * _READ_CHUNK:(k + 1) * _READ_CHUNK]))
        if first_bad >= len(data):
            dec.decode(b"", final=True)  # the empty read at the end of the file flushes the decoder
    except UnicodeDecodeError as exc:
        return f"{type(exc).__name__}: {exc}"
    return "UnicodeDecodeError: 'utf-8' codec can't decode bytes: invalid data"


def reading_failure(info, index, data, stop_row: int) -> Optional[str]:
    """What the reference's reader would raise before it has read data row `stop_row` (it reads
    that row too before it stops: ops/csv_shard.py:19-24), as "Type: message", or None.

    Two things can go wrong while it reads: the strict UTF-8 decoder rejects a byte in one of the
    8 KiB reads it has pulled so far, or a field outgrows csv.field_size_limit()."""
    n = int(info.n_bytes)
    rows = int(info.rows)
    if stop_row >= rows:
        needed, to_eof = n - 1, True
    else:
        _, end = index.row_spans(stop_row, 1)
        needed, to_eof = int(end[0]), False
        if needed >= n:
            needed, to_eof = n - 1, True
        elif data[needed:needed + 1] == b"\r" and needed + 1 < n:
            needed += 1

    limit_at = None  # last byte of the line whose parsing trips the field limit
    if info.header_field_too_long:
        limit_at = _line_end_after(data, int(info.header_long_offset))
    elif 0 <= int(info.long_field_row) <= stop_row:
        limit_at = _line_end_after(data, int(info.long_field_offset))

    bad = int(info.first_invalid_utf8)
    if bad >= 0:
        horizon = limit_at if limit_at is not None else needed
        hit = bad // _READ_CHUNK <= horizon // _READ_CHUNK if bad < n else (to_eof and limit_at is None)
        if hit:
            return _decode_error_text(data, bad)
    if limit_at is not None:
        return "Error: field larger than field limit (131072)"
    return None


def build_rows(names: Optional[List[str]], blob, field_end, row_field_end) -> List[Dict[Any, Any]]:
    """DictReader's row dicts from the unquoted field content: names zipped with values, surplus
    values under the key None, missing values None (csv.DictReader.__next__)."""
    raw = blob.tobytes()
    ends = field_end.tolist()
    if raw.isascii():
        text = raw.decode("ascii")
        cut = lambda a, b: text[a:b]
    else:
        cut = lambda a, b: raw[a:b].decode("utf-8")
    names = names or []
    k = len(names)
    rows: List[Dict[Any, Any]] = []
    f = 0
    pos = 0
    for stop in row_field_end.tolist():
        values = []
        for j in range(f, stop):
            values.append(cut(pos, ends[j]))
            pos = ends[j]
        f = stop
        row: Dict[Any, Any] = dict(zip(names, values))
        if k < len(values):
            row[None] = values[k:]
        elif k > len(values):
            for name in names[len(values):]:
                row[name] = None
        rows.append(row)
    return rows


def open_index(source_uri: str):
    from ._cuda_runtime import get_csv_index, get_tokenizer_engine

    return get_csv_index(source_uri, get_tokenizer_engine())


@register_op("read_csv_shard")
@register_op("csv_shard")
def op_read_csv_shard(task_or_payload: Optional[Dict[str, Any]]) -> Dict[str, Any]:
    if task_or_payload is None:
        return _fail("missing payload")
    if not isinstance(task_or_payload, dict):
        return _fail("payload must be a dict")
    payload = task_or_payload["payload"] if "payload" in task_or_payload else task_or_payload
    if not isinstance(payload, dict):
        return _fail("payload must be a dict")

    source_uri = payload.get("source_uri")
    if not source_uri or not isinstance(source_uri, str):
        return _fail("payload.source_uri (string) is required")
    try:
        start_row = int(payload.get("start_row", 0))
        shard_size = int(payload.get("shard_size", 100))
    except Exception:
        return _fail("start_row and shard_size must be integers")
    if start_row < 0:
        return _fail("start_row must be >= 0")
    if shard_size <= 0:
        return _fail("shard_size must be > 0")
    mode = payload.get("mode", "rows")
    if mode not in ("rows", "count"):
        return _fail("mode must be 'rows' or 'count'")
    if not os.path.exists(source_uri):
        return _fail(f"file not found: {source_uri}")

    try:
        index, data = open_index(source_uri)
        info = index.info
        problem = reading_failure(info, index, data, start_row + shard_size)
        if problem:
            return _fail(f"failed reading csv: {problem}")
        first = min(start_row, index.rows)
        count = min(shard_size, index.rows - first)
        rows = None
        if mode == "rows":
            blob, field_end, row_field_end = index.rows_read(first, count)
            rows = build_rows(index.fieldnames(), blob, field_end, row_field_end)
    except Exception as exc:
        return _fail(f"failed reading csv: {type(exc).__name__}: {exc}")

    out = {
        "ok": True,
        "dataset_id": payload.get("dataset_id", "unknown_dataset"),
        "mode": mode,
        "start_row": start_row,
        "end_row": start_row + count,
        "row_count": count,
    }
    if rows is not None:
        out["rows"] = rows
    return out
