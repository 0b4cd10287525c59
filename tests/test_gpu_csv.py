"""CSV path on the GPU: the record index, `read_csv_shard` and the csv -> classify chain.

Integer / byte work, so everything is compared bit for bit: the device index against the oracle's
byte-level restatement (oracle/csv_oracle.py), the op against the reference's recorded results
(tests/golden/csv_golden.json) and against the oracle op (the reference's algorithm on CPython's
csv module) on larger generated tables."""
import csv
import hashlib
import io
import json
import os

import numpy as np
import pytest

import csv_oracle
from csv_cases import _table, all_cases

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "csv_golden.json")


def canonical(result, path):
    out = {}
    for k, v in result.items():
        if k == "rows":
            out[k] = [[[key, val] for key, val in row.items()] for row in v]
        elif isinstance(v, str):
            out[k] = v.replace(path, "<FILE>")
        else:
            out[k] = v
    return json.loads(json.dumps(out))


def matches(expected, got):
    if isinstance(expected, dict) and set(expected) == {"sha256", "chars"}:
        text = json.dumps(got, sort_keys=True, ensure_ascii=True, separators=(",", ":"))
        return hashlib.sha256(text.encode()).hexdigest() == expected["sha256"]
    return expected == got


@pytest.fixture(scope="module")
def plain_engine():
    from ops import _cuda_runtime as rt
    eng = rt.Engine(None, None, max_batch=1, max_len=128)
    yield eng
    eng.close()


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN, encoding="utf-8") as f:
        return json.load(f)


def _check_index(eng, name, data):
    from ops import _cuda_runtime as rt
    want = csv_oracle.index(data)
    idx = rt.CsvIndex(eng, data)
    try:
        info = idx.info
        for key in ("n_bytes", "has_header", "header_fields", "data_offset", "rows", "first_invalid_utf8",
                    "header_field_too_long", "header_long_offset", "long_field_row", "long_field_offset"):
            assert int(getattr(info, key)) == int(want[key]), (name, key)
        if want["first_invalid_utf8"] < 0:
            names = idx.fieldnames()
            assert names == (None if want["header"] is None else [b.decode() for b in want["header"]]), name
        n = want["rows"]
        a, e = idx.row_spans(0, n)
        assert (a == want["row_start"]).all() and (e == want["row_end"]).all(), name
        for start, count, col in ((0, n, -1), (n // 3, n - n // 3, -1), (0, n, 1), (0, n, 0), (0, n, 7),
                                  (n // 2, min(5, n - n // 2), 2)):
            blob, fend, rend = idx.rows_read(start, count, col)
            wb, wf, wr = csv_oracle.rows_content(want, start, count, col)
            assert blob.tobytes() == wb.tobytes(), (name, start, count, col)
            assert (fend == wf).all() and (rend == wr).all(), (name, start, count, col)
    finally:
        idx.close()


def test_index_matches_oracle_on_every_case(plain_engine):
    for name, (data, _) in all_cases().items():
        _check_index(plain_engine, name, data)


def test_op_matches_reference_golden(golden, tmp_path, monkeypatch):
    from ops import _cuda_runtime as rt, get_op
    rt.reset_cache()
    op = get_op("read_csv_shard")
    for name, (data, queries) in all_cases().items():
        path = os.path.join(str(tmp_path), name + ".csv")
        with open(path, "wb") as f:
            f.write(data)
        for entry, (start_row, shard_size, mode) in zip(golden[name]["queries"], queries):
            res = op({"source_uri": path, "start_row": start_row, "shard_size": shard_size, "mode": mode,
                      "dataset_id": name})
            assert matches(entry["result"], canonical(res, path)), (name, entry["query"], res.get("error"))
    rt.reset_cache()


def test_large_table_against_oracle_op(tmp_path):
    """A 60K-row id,text,source table with quoting of every kind, CRLF line ends: shards cut from
    the device index equal the reference algorithm's (csv.DictReader) for rows and counts."""
    from ops import _cuda_runtime as rt, get_op
    rt.reset_cache()
    op = get_op("csv_shard")
    data = _table(60000, 77, "\r\n")
    path = os.path.join(str(tmp_path), "big.csv")
    with open(path, "wb") as f:
        f.write(data)
    for start_row, shard_size, mode in ((0, 100, "rows"), (59990, 100, "rows"), (12345, 5000, "rows"),
                                        (0, 10 ** 6, "count"), (30000, 1, "rows"), (70000, 5, "rows")):
        payload = {"source_uri": path, "start_row": start_row, "shard_size": shard_size, "mode": mode}
        assert op(dict(payload)) == csv_oracle.read_csv_shard(dict(payload)), (start_row, shard_size, mode)
    rt.reset_cache()


def test_multi_megabyte_scan_properties(plain_engine):
    """~40 MB: row count, spans and a checksum of the text column against a streaming csv.reader."""
    from ops import _cuda_runtime as rt
    part = _table(20000, 5)
    body = part[part.index(b"\n") + 1:]
    data = b"id,text,source\n" + body * 24
    idx = rt.CsvIndex(plain_engine, data)
    try:
        want_rows = 0
        want_sha = hashlib.sha256()
        for i, rec in enumerate(csv.reader(io.StringIO(data.decode("utf-8"), newline=""))):
            if i == 0 or not rec:
                continue
            want_rows += 1
            want_sha.update(rec[1].encode("utf-8"))
        assert idx.rows == want_rows
        blob, fend, rend = idx.rows_read(0, idx.rows, 1)
        assert hashlib.sha256(blob.tobytes()).hexdigest() == want_sha.hexdigest()
        assert (rend == np.arange(1, idx.rows + 1)).all()
        a, e = idx.row_spans(0, idx.rows)
        assert (a[1:] > e[:-1]).all() and (e >= a).all()
    finally:
        idx.close()


def test_csv_chain_equals_classifying_the_column(tiny_model, vocab_path, tmp_path, monkeypatch):
    """map_classify_tpu with source_uri == map_classify_tpu with items = that column read by the
    reference algorithm (missing cells read as "", as map_tokenize does: ops/map_tokenize.py:38)."""
    from ops import _cuda_runtime as rt, get_op
    rt.reset_cache()
    monkeypatch.setenv("CUDA_MODEL_PATH", tiny_model)
    monkeypatch.setenv("CUDA_VOCAB_PATH", vocab_path)
    data = _table(3000, 11) + b"7777\n" + b'8888,"last row, no line end",z'
    path = os.path.join(str(tmp_path), "chain.csv")
    with open(path, "wb") as f:
        f.write(data)
    classify = get_op("map_classify_tpu")
    for start_row, shard_size in ((0, None), (100, 1000), (2990, 50)):
        payload = {"source_uri": path, "start_row": start_row, "return_results": True, "return_logits": True,
                   "batch": 64, "allow_fallback": False}
        if shard_size:
            payload["shard_size"] = shard_size
        got = classify(dict(payload))
        rows = csv_oracle.read_csv_shard({"source_uri": path, "start_row": start_row,
                                          "shard_size": shard_size or 10 ** 9})["rows"]
        items = [r.get("text") for r in rows]
        want = classify({"items": items, "return_results": True, "return_logits": True, "batch": 64,
                         "allow_fallback": False})
        assert got["row_count"] == len(items) == got["items_count"]
        assert got["end_row"] == start_row + len(items)
        for key in ("labels", "label_histogram", "results", "logits", "topk"):
            assert got[key] == want[key], (start_row, key)
        assert got["metrics"]["tokens"] == want["metrics"]["tokens"]
    # error envelope: unknown column, unreadable file
    bad = classify({"source_uri": path, "text_column": "nope"})
    assert bad["fallback"] == "cpu" and "nope" in bad["reason"]
    with open(path, "ab") as f:
        f.write(b"\n1,\xff,bad\n")
    bad = classify({"source_uri": path})
    assert bad["fallback"] == "cpu" and "UnicodeDecodeError" in bad["reason"]
    rt.reset_cache()
