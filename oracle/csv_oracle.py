"""CPU oracle for the CSV path — TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), the
cpu_baseline / reference legs of bench.py). The product path never imports it.

Two layers:

1. `read_csv_shard(payload)` restates the reference op (ops/csv_shard.py:29-103 with its helper
   `_read_csv_shard`, :9-26) on top of the same third-party code the reference calls — CPython's
   `csv.DictReader` (stdlib `csv` / `Modules/_csv.c`, CPython 3.12.3 in this image) over
   `open(path, newline="", encoding="utf-8")`. It is pinned by tests/golden/csv_golden.json, which
   tests/golden/make_csv_golden.py records from the reference module itself.

2. `index(data)` / `rows_content(...)` restate, byte by byte in plain Python, what the CUDA scan
   must produce (row spans, unquoted field content, the two failure positions). It follows the
   reader's state machine in `Modules/_csv.c` (`parse_process_char`: START_RECORD, START_FIELD,
   IN_FIELD, IN_QUOTED_FIELD, QUOTE_IN_QUOTED_FIELD, EAT_CRNL with the default dialect) and
   `Reader_iternext` (a record ends when a physical line ends outside quotes; at end of input an
   open field or quoted field is still delivered), and is pinned against `csv.reader` itself in
   tests/test_csv_oracle.py. Small inputs only: it is a per-byte Python loop.
"""
from __future__ import annotations

import csv
import os
from typing import Any, Dict, List, Optional

import numpy as np

FIELD_LIMIT = 131072  # csv.field_size_limit() default


# ---- layer 1: the reference op on stdlib csv --------------------------------------------------------

def _slice_rows(path: str, start_row: int, shard_size: int) -> List[Dict[Any, Any]]:
    """ops/csv_shard.py:9-26: iterate DictReader, keep rows start_row <= idx < start_row + shard_size."""
    stop = start_row + shard_size
    kept: List[Dict[Any, Any]] = []
    with open(path, newline="", encoding="utf-8") as handle:
        for idx, row in enumerate(csv.DictReader(handle)):
            if idx >= stop:
                break
            if idx >= start_row:
                kept.append(row)
    return kept


def read_csv_shard(task_or_payload: Optional[Dict[str, Any]]) -> Dict[str, Any]:
    """ops/csv_shard.py:30-103, envelope and checks in the reference's order."""
    def no(msg):
        return {"ok": False, "error": "read_csv_shard: " + msg}

    if task_or_payload is None:
        return no("missing payload")
    if not isinstance(task_or_payload, dict):
        return no("payload must be a dict")
    payload = task_or_payload.get("payload") if "payload" in task_or_payload else task_or_payload
    if not isinstance(payload, dict):
        return no("payload must be a dict")
    uri = payload.get("source_uri")
    if not uri or not isinstance(uri, str):
        return no("payload.source_uri (string) is required")
    try:
        start_row = int(payload.get("start_row", 0))
        shard_size = int(payload.get("shard_size", 100))
    except Exception:
        return no("start_row and shard_size must be integers")
    if start_row < 0:
        return no("start_row must be >= 0")
    if shard_size <= 0:
        return no("shard_size must be > 0")
    mode = payload.get("mode", "rows")
    if mode not in ("rows", "count"):
        return no("mode must be 'rows' or 'count'")
    if not os.path.exists(uri):
        return no(f"file not found: {uri}")
    try:
        rows = _slice_rows(uri, start_row, shard_size)
    except Exception as exc:
        return no(f"failed reading csv: {type(exc).__name__}: {exc}")
    out = {"ok": True, "dataset_id": payload.get("dataset_id", "unknown_dataset"), "mode": mode,
           "start_row": start_row, "end_row": start_row + len(rows), "row_count": len(rows)}
    if mode == "rows":
        out["rows"] = rows
    return out


# ---- layer 2: what the device index must hold ---------------------------------------------------------

def _records(data: bytes):
    """Yield (start, end, fields, overflow) for every record of `data` in file order, empty records
    included (fields == []): `start` is the first byte, `end` the offset of the line end closing it
    (len(data) without one), `fields` the unquoted field contents as bytes, `overflow` the offset
    of the first code point that pushes a field past FIELD_LIMIT (or -1)."""
    n = len(data)
    pos = 0
    while pos < n:
        start = pos
        state = "start_record"
        fields: List[bytes] = []
        cur = bytearray()
        cur_chars = 0
        overflow = -1

        def add(b, at):
            nonlocal cur_chars, overflow
            if b & 0xC0 != 0x80:
                if cur_chars >= FIELD_LIMIT and overflow < 0:
                    overflow = at
                cur_chars += 1
            cur.append(b)

        def save():
            nonlocal cur, cur_chars
            fields.append(bytes(cur))
            cur = bytearray()
            cur_chars = 0

        end = None
        while pos < n:
            b = data[pos]
            eol = b in (0x0A, 0x0D)
            if state == "start_record":
                if eol:
                    end = pos
                    break
                state = "start_field"
            if state == "start_field":
                if eol:
                    save()
                    end = pos
                    break
                if b == 0x22:
                    state = "quoted"
                elif b == 0x2C:
                    save()
                else:
                    add(b, pos)
                    state = "field"
            elif state == "field":
                if eol:
                    save()
                    end = pos
                    break
                if b == 0x2C:
                    save()
                    state = "start_field"
                else:
                    add(b, pos)
            elif state == "quoted":
                if b == 0x22:
                    state = "quote_in_quoted"
                else:
                    add(b, pos)
            else:  # quote_in_quoted
                if b == 0x22:
                    add(b, pos)
                    state = "quoted"
                elif b == 0x2C:
                    save()
                    state = "start_field"
                elif eol:
                    save()
                    end = pos
                    break
                else:
                    add(b, pos)
                    state = "field"
            pos += 1
        if end is None:
            # input ended: Reader_iternext still delivers an open record
            end = n
            if state != "start_record":
                save()
            yield start, end, fields, overflow
            return
        yield start, end, fields, overflow
        # step over the line end: "\r\n" is one
        pos = end + 1
        if data[end] == 0x0D and pos < n and data[pos] == 0x0A:
            pos += 1


def first_invalid_utf8(data: bytes) -> int:
    """Offset at which a strict UTF-8 decoder holds the evidence of an error (the byte that cannot be
    there; len(data) when the input ends inside a sequence), or -1."""
    n = len(data)
    i = 0
    while i < n:
        b = data[i]
        if b < 0x80:
            i += 1
            continue
        if b < 0xC2 or b > 0xF4:
            return i
        need = 2 if b < 0xE0 else 3 if b < 0xF0 else 4
        for j in range(1, need):
            if i + j >= n:
                return n
            c = data[i + j]
            lo, hi = 0x80, 0xBF
            if j == 1:
                if b == 0xE0:
                    lo = 0xA0
                elif b == 0xED:
                    hi = 0x9F
                elif b == 0xF0:
                    lo = 0x90
                elif b == 0xF4:
                    hi = 0x8F
            if not lo <= c <= hi:
                return i + j
        i += need
    return -1


def index(data: bytes) -> Dict[str, Any]:
    """The fields of b2_csv_info plus the row spans and the header names (as bytes)."""
    recs = list(_records(data))
    out: Dict[str, Any] = {"n_bytes": len(data), "has_header": 0, "header": None, "header_fields": 0,
                           "data_offset": len(data), "header_field_too_long": 0, "header_long_offset": -1}
    body = recs
    if recs:
        start, end, fields, overflow = recs[0]
        out["has_header"] = 1
        out["header"] = fields
        out["header_fields"] = len(fields)
        out["header_field_too_long"] = int(overflow >= 0)
        out["header_long_offset"] = overflow
        off = end + 1 if end < len(data) else len(data)
        if end < len(data) and data[end] == 0x0D and off < len(data) and data[off] == 0x0A:
            off += 1
        out["data_offset"] = off
        body = recs[1:]
    rows = [r for r in body if r[2]]  # DictReader drops records without fields
    out["rows"] = len(rows)
    out["row_start"] = np.array([r[0] for r in rows], dtype=np.int64)
    out["row_end"] = np.array([r[1] for r in rows], dtype=np.int64)
    out["row_fields"] = [r[2] for r in rows]
    long_rows = [(i, r[3]) for i, r in enumerate(rows) if r[3] >= 0]
    out["long_field_row"], out["long_field_offset"] = long_rows[0] if long_rows else (-1, -1)
    out["first_invalid_utf8"] = first_invalid_utf8(data)
    return out


def rows_content(idx: Dict[str, Any], start_row: int, n_rows: int, column: int = -1):
    """(content bytes, field_end, row_field_end) as b2_csv_rows_read returns them."""
    blob = bytearray()
    field_end: List[int] = []
    row_field_end: List[int] = []
    for fields in idx["row_fields"][start_row:start_row + n_rows]:
        if column >= 0:
            fields = [fields[column] if column < len(fields) else b""]
        for f in fields:
            blob += f
            field_end.append(len(blob))
        row_field_end.append(len(field_end))
    return (np.frombuffer(bytes(blob), dtype=np.uint8), np.array(field_end, dtype=np.int64),
            np.array(row_field_end, dtype=np.int64))


class OracleIndex:
    """Duck-typed stand-in for ops._cuda_runtime.CsvIndex, so that the op's host logic
    (reading_failure, build_rows) can be exercised without a GPU."""

    class _Info:
        pass

    def __init__(self, data: bytes):
        self._idx = index(data)
        self.info = self._Info()
        for k in ("n_bytes", "has_header", "header_fields", "data_offset", "rows", "first_invalid_utf8",
                  "header_field_too_long", "header_long_offset", "long_field_row", "long_field_offset"):
            setattr(self.info, k, self._idx[k])
        self.rows = self._idx["rows"]
        self.n_bytes = len(data)

    def fieldnames(self):
        h = self._idx["header"]
        return None if h is None else [b.decode("utf-8") for b in h]

    def row_spans(self, start_row, n_rows):
        return (self._idx["row_start"][start_row:start_row + n_rows],
                self._idx["row_end"][start_row:start_row + n_rows])

    def rows_read(self, start_row, n_rows, column=-1):
        return rows_content(self._idx, start_row, n_rows, column)
