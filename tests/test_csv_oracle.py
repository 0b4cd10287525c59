"""CPU tests of the CSV path: the oracle is pinned against the reference's recorded outputs
(tests/golden/csv_golden.json) and against CPython's csv reader, and the op's host logic (envelope,
failure rules, row dict assembly) is run on top of the oracle's index in place of the GPU's."""
import csv
import hashlib
import io
import json
import os

import pytest

import csv_oracle
from csv_cases import all_cases, small_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "csv_golden.json")
PLACEHOLDER = "<FILE>"


def canonical(result, path):
    out = {}
    for k, v in result.items():
        if k == "rows":
            out[k] = [[[key, val] for key, val in row.items()] for row in v]
        elif isinstance(v, str):
            out[k] = v.replace(path, PLACEHOLDER)
        else:
            out[k] = v
    return out


def matches(expected, got_canonical):
    got = json.loads(json.dumps(got_canonical))  # tuples -> lists, None keys already as pairs
    if isinstance(expected, dict) and set(expected) == {"sha256", "chars"}:
        text = json.dumps(got, sort_keys=True, ensure_ascii=True, separators=(",", ":"))
        return hashlib.sha256(text.encode()).hexdigest() == expected["sha256"]
    return expected == got


@pytest.fixture(scope="module")
def golden():
    with open(GOLDEN, encoding="utf-8") as f:
        return json.load(f)


@pytest.fixture(scope="module")
def case_files(tmp_path_factory):
    root = tmp_path_factory.mktemp("csvcases")
    paths = {}
    for name, (data, queries) in all_cases().items():
        p = os.path.join(str(root), name + ".csv")
        with open(p, "wb") as f:
            f.write(data)
        paths[name] = (p, data, queries)
    return paths


def test_cases_match_the_recorded_inputs(golden):
    for name, (data, _) in all_cases().items():
        assert golden[name]["sha256"] == hashlib.sha256(data).hexdigest(), name


def test_oracle_op_matches_reference_golden(golden, case_files):
    """oracle.read_csv_shard == what the reference's op_read_csv_shard returned (recorded)."""
    for name, (path, _, queries) in case_files.items():
        for entry, (start_row, shard_size, mode) in zip(golden[name]["queries"], queries):
            assert entry["query"] == [start_row, shard_size, mode]
            res = csv_oracle.read_csv_shard({"source_uri": path, "start_row": start_row,
                                             "shard_size": shard_size, "mode": mode, "dataset_id": name})
            assert matches(entry["result"], canonical(res, path)), (name, entry["query"])


def test_oracle_index_matches_csv_reader():
    """The byte-level restatement yields exactly the records csv.reader yields."""
    for name, data in small_cases().items():
        recs = list(csv.reader(io.StringIO(data.decode("utf-8"), newline="")))
        idx = csv_oracle.index(data)
        if not recs:
            assert idx["has_header"] == 0 and idx["rows"] == 0, name
            continue
        assert idx["has_header"] == 1, name
        assert [b.decode() for b in idx["header"]] == recs[0], name
        want = [r for r in recs[1:] if r]
        got = [[b.decode() for b in fields] for fields in idx["row_fields"]]
        assert got == want, name
        # spans: every row starts on its first byte and ends on a line end (or at the end of data)
        for a, e in zip(idx["row_start"].tolist(), idx["row_end"].tolist()):
            assert data[a] not in b"\r\n", name
            assert e == len(data) or data[e] in b"\r\n", name


def test_first_invalid_utf8_matches_python_decoder():
    samples = [b"", b"plain", "é東😀".encode(), b"\xff", b"a\x80", b"\xc3", b"\xc3\x28", b"ab\xe6\x9d",
               b"\xe0\x80\xaf", b"\xed\xa0\x80", b"\xf4\x90\x80\x80", b"\xf0\x9f\x98", b"\xf5\x80",
               b"x\xe2\x28\xa1", b"\xc0\xaf", b"\xc1\xbf", b"ok\xf0\x9f\x98\x80ok\x80"]
    for s in samples:
        got = csv_oracle.first_invalid_utf8(s)
        try:
            s.decode("utf-8")
            assert got == -1, s
        except UnicodeDecodeError as exc:
            # Python reports [start, end) of the bad sequence; the evidence sits at `end` unless the
            # start byte itself is illegal (then end == start + 1) or the data ran out
            if exc.reason == "unexpected end of data":
                assert got == len(s), s
            elif exc.reason == "invalid start byte":
                assert got == exc.start, s
            else:
                assert got == exc.end, (s, exc.start, exc.end, got)


def test_op_host_logic_on_oracle_index(golden, case_files, monkeypatch):
    """agent-tpu_b200/ops/csv_shard.py with the oracle's index standing in for the GPU's: envelope,
    failure rules and row assembly give the reference's recorded results for every case."""
    from ops import csv_shard

    for name, (path, data, queries) in case_files.items():
        oi = csv_oracle.OracleIndex(data)
        monkeypatch.setattr(csv_shard, "open_index", lambda uri, oi=oi, data=data: (oi, data))
        for entry, (start_row, shard_size, mode) in zip(golden[name]["queries"], queries):
            res = csv_shard.op_read_csv_shard({"source_uri": path, "start_row": start_row,
                                               "shard_size": shard_size, "mode": mode, "dataset_id": name})
            assert matches(entry["result"], canonical(res, path)), (name, entry["query"], res.get("error"))


def test_op_envelope_errors_match_reference(golden, tmp_path):
    from ops import csv_shard, get_op

    assert get_op("read_csv_shard") is csv_shard.op_read_csv_shard
    assert get_op("csv_shard") is csv_shard.op_read_csv_shard
    missing = os.path.join(str(tmp_path), "does_not_exist.csv")
    args = {"none": None, "not_dict": [1, 2], "payload_none": {"payload": None}, "payload_list": {"payload": [1]},
            "no_uri": {}, "uri_not_str": {"source_uri": 5}, "bad_ints": {"source_uri": "x", "start_row": "abc"},
            "neg_start": {"source_uri": "x", "start_row": -1}, "zero_size": {"source_uri": "x", "shard_size": 0},
            "bad_mode": {"source_uri": "x", "mode": "both"}, "missing_file": {"source_uri": missing},
            "task_wrapped_missing": {"id": "t1", "payload": {"source_uri": missing, "mode": "count"}}}
    for entry in golden["__envelope__"]:
        if entry["label"] not in args:
            continue
        for fn in (csv_shard.op_read_csv_shard, csv_oracle.read_csv_shard):
            res = fn(args[entry["label"]])
            assert canonical(res, missing) == entry["result"], (entry["label"], fn.__module__)
