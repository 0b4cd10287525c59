#!/usr/bin/env python3
"""Generate tests/golden/csv_golden.json by IMPORTING the reference's ops/csv_shard.py (read-only)
and recording what `op_read_csv_shard` returns for every case and query of tests/csv_cases.py.

Run in the build container only (the GPU box has no /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_csv_golden.py

Results are stored canonically (see `canonical` below: row dicts as [key, value] pair lists, since
DictReader uses the key None for surplus fields); results above 4 KB as their SHA-256.
"""
import hashlib
import json
import os
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

from csv_cases import all_cases  # noqa: E402
from ops.csv_shard import op_read_csv_shard  # noqa: E402  (the reference)

PLACEHOLDER = "<FILE>"


def canonical(result, path):
    """JSON-able, order-stable form of an op result; the temp file's path is masked."""
    out = {}
    for k, v in result.items():
        if k == "rows":
            out[k] = [[[key, val] for key, val in row.items()] for row in v]
        elif isinstance(v, str):
            out[k] = v.replace(path, PLACEHOLDER)
        else:
            out[k] = v
    return out


def digest(obj):
    text = json.dumps(obj, sort_keys=True, ensure_ascii=True, separators=(",", ":"))
    if len(text) <= 4096:
        return obj
    return {"sha256": hashlib.sha256(text.encode()).hexdigest(), "chars": len(text)}


def main():
    golden = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, (data, queries) in all_cases().items():
            path = os.path.join(tmp, name + ".csv")
            with open(path, "wb") as f:
                f.write(data)
            entries = []
            for start_row, shard_size, mode in queries:
                payload = {"source_uri": path, "start_row": start_row, "shard_size": shard_size, "mode": mode,
                           "dataset_id": name}
                entries.append({"query": [start_row, shard_size, mode],
                                "result": digest(canonical(op_read_csv_shard(payload), path))})
            golden[name] = {"sha256": hashlib.sha256(data).hexdigest(), "bytes": len(data), "queries": entries}
        # envelope checks that need no file
        env = []
        missing = os.path.join(tmp, "does_not_exist.csv")
        for label, arg in [("none", None), ("not_dict", [1, 2]), ("payload_none", {"payload": None}),
                           ("payload_list", {"payload": [1]}), ("no_uri", {}), ("uri_not_str", {"source_uri": 5}),
                           ("bad_ints", {"source_uri": "x", "start_row": "abc"}),
                           ("neg_start", {"source_uri": "x", "start_row": -1}),
                           ("zero_size", {"source_uri": "x", "shard_size": 0}),
                           ("bad_mode", {"source_uri": "x", "mode": "both"}),
                           ("missing_file", {"source_uri": missing}),
                           ("task_wrapped_missing", {"id": "t1", "payload": {"source_uri": missing, "mode": "count"}}),
                           ("is_directory", {"source_uri": tmp})]:
            res = op_read_csv_shard(arg)
            env.append({"label": label, "result": canonical(res, missing if "missing" in label else tmp)})
        golden["__envelope__"] = env
    with open(os.path.join(HERE, "csv_golden.json"), "w", encoding="utf-8") as f:
        json.dump(golden, f, ensure_ascii=True, sort_keys=True, indent=0, separators=(",", ":"))
    print("cases:", len(golden) - 1, "bytes:", os.path.getsize(os.path.join(HERE, "csv_golden.json")))


if __name__ == "__main__":
    main()
