#!/usr/bin/env python3
"""GPU box: BASELINE config 5 on one GPU — a synthetic `id,text,source` CSV is indexed on the device
(b2_csv_open), counted / sliced through `read_csv_shard`, and classified through the chained
`map_classify_tpu` payload (source_uri). Prints one JSON line per measurement.

    python tests/tools/gpu_csv_bench.py [rows=2000000] [classify_rows=262144]

The CPU number beside the index is the reference's algorithm (csv.DictReader, oracle/csv_oracle.py)
counting the same file on one host core (it is a single-threaded C loop)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import model_file as mf  # noqa: E402
import synth  # noqa: E402
from ops import _cuda_runtime as rt, get_op  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
classify_rows = int(sys.argv[2]) if len(sys.argv) > 2 else 262_144
tmp = "/tmp/b2csv"
os.makedirs(tmp, exist_ok=True)
vocab = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
model = mf.write_model_file(os.path.join(tmp, "base.b2m"), mf.BERT_BASE)
os.environ["CUDA_MODEL_PATH"] = model
os.environ["CUDA_VOCAB_PATH"] = vocab

# ---- the file: sentence-case text with commas, so every text cell is quoted; 1 in 16 carries a quote
pool = synth.SentenceGenerator().sentences(16384)
cells = []
for i, s in enumerate(pool):
    if i % 16 == 0:
        s = s.replace(" ", ' "q" ', 1)
    cells.append('"' + s.replace('"', '""') + '"' if any(ch in s for ch in ',"\n') else s)
path = os.path.join(tmp, f"synthetic_{rows}.csv")
t = time.time()
with open(path, "w", encoding="utf-8", newline="") as f:
    f.write("id,text,source\n")
    block = []
    for i in range(rows):
        block.append(f"{i},{cells[i % len(cells)]},src{i % 10}\n")
        if len(block) == 65536:
            f.write("".join(block))
            block = []
    f.write("".join(block))
size = os.path.getsize(path)
print(json.dumps({"file": path, "rows": rows, "bytes": size, "write_s": round(time.time() - t, 2)}), flush=True)

# ---- index: copy + scan, three times (first includes page-cache and allocator warm-up)
eng = rt.get_cuda_handle(model, vocab, max_len=128, max_batch=1024).interpreter
import mmap
with open(path, "rb") as f:
    data = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
for it in range(3):
    t = time.time()
    idx = rt.CsvIndex(eng, data)
    wall = time.time() - t
    info = idx.info
    print(json.dumps({"what": "b2_csv_open", "run": it, "rows": idx.rows, "bytes": size,
                      "wall_ms": round(wall * 1e3, 2), "gpu_ms": round(info.gpu_ms, 3),
                      "index_ms": round(info.index_ms, 3), "launches": int(info.kernel_launches),
                      "index_GBps_file_bytes": round(size / info.index_ms / 1e6, 1),
                      "e2e_rows_per_s": round(idx.rows / wall)}), flush=True)
    if it < 2:
        idx.close()
assert idx.rows == rows

# ---- field extraction on the device: the text column of every row
t = time.time()
blob, fend, rend = idx.rows_read(0, idx.rows, 1)
print(json.dumps({"what": "rows_read(text column)", "rows": idx.rows, "content_bytes": int(blob.size),
                  "wall_ms": round((time.time() - t) * 1e3, 2)}), flush=True)
idx.close()
data.close()
if classify_rows == 0:
    sys.exit(0)  # index-only run (the one to put under ncu)

# ---- the ops
op = get_op("read_csv_shard")
for label, payload in (("count all (first call: open + index)", {"mode": "count", "shard_size": 10 ** 9}),
                       ("count all (index resident)", {"mode": "count", "shard_size": 10 ** 9}),
                       ("rows shard of 100 at the end", {"start_row": rows - 100, "shard_size": 100}),
                       ("rows shard of 100000", {"start_row": rows // 2, "shard_size": 100000})):
    t = time.time()
    res = op({"source_uri": path, **payload})
    wall = time.time() - t
    assert res["ok"], res
    print(json.dumps({"what": "read_csv_shard " + label, "row_count": res["row_count"],
                      "wall_ms": round(wall * 1e3, 2)}), flush=True)

import csv_oracle  # noqa: E402
t = time.time()
ref = csv_oracle.read_csv_shard({"source_uri": path, "mode": "count", "shard_size": 10 ** 9})
wall = time.time() - t
assert ref["row_count"] == rows
print(json.dumps({"what": "cpu_baseline csv.DictReader count all (reference algorithm, 1 core)",
                  "wall_ms": round(wall * 1e3, 1), "rows_per_s": round(rows / wall),
                  "MBps": round(size / wall / 1e6, 1)}), flush=True)
t = time.time()
ref = csv_oracle.read_csv_shard({"source_uri": path, "start_row": rows - 100, "shard_size": 100})
print(json.dumps({"what": "cpu_baseline rows shard of 100 at the end", "wall_ms": round((time.time() - t) * 1e3, 1)}),
      flush=True)

# ---- the chain
classify = get_op("map_classify_tpu")
n = min(classify_rows, rows)
for it in range(2):
    t = time.time()
    res = classify({"source_uri": path, "start_row": 0, "shard_size": n, "allow_fallback": False})
    wall = time.time() - t
    m = res["metrics"]
    print(json.dumps({"what": "chain csv -> tokenize -> classify", "run": it, "rows": res["row_count"],
                      "wall_ms": round(wall * 1e3, 1), "items_per_s_wall": round(n / wall),
                      "gpu_ms": round(m["gpu_ms_total"], 1), "items_per_s_gpu": round(n / m["gpu_ms_total"] * 1e3),
                      "tokens": m["tokens"], "launches": m["kernel_launches"], "csv": m["csv"],
                      "hist_sum": sum(res["label_histogram"])}), flush=True)
