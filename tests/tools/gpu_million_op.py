#!/usr/bin/env python3
"""GPU box: BASELINE config 3 as one literal job — ops.get_op("map_classify_tpu")({"items": [1,000,000
strings]}) called ONCE from ONE process, over every GPU of the box (EngineGroup), with the three result
forms of SURVEY.md section 8(f)-2: JSON lists (default), result_format "packed", result_uri (.npz file).
Prints one JSON line per run: wall time of the op call, the device time inside it, the size of the result
as the agent would post it (json.dumps), and that the three forms carry the same labels.

    python tests/tools/gpu_million_op.py [items=1000000]
"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
import model_file as mf  # noqa: E402
import synth  # noqa: E402
import ops  # noqa: E402
from ops import _cuda_runtime as rt  # noqa: E402
from ops import map_classify_tpu as op  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
tmp = tempfile.mkdtemp(prefix="b2million")
os.environ["CUDA_VOCAB_PATH"] = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
os.environ["CUDA_MODEL_PATH"] = mf.write_model_file(os.path.join(tmp, "base.b2m"), mf.BERT_BASE)
pool = synth.SentenceGenerator().sentences(4096)
items = (pool * (n // len(pool) + 1))[:n]
run = ops.get_op("map_classify_tpu")
run({"items": items[:2048], "allow_fallback": False})                      # loads the handle(s)
devices = rt.job_devices()
labels = {}
for name, extra in (("json lists", {}), ("packed", {"result_format": "packed"}),
                    ("result_uri", {"result_uri": os.path.join(tmp, "result")})):
    t0 = time.perf_counter()
    r = run({"items": items, "topk": 5, "return_results": name == "json lists", "allow_fallback": False, **extra})
    wall = time.perf_counter() - t0
    t1 = time.perf_counter()
    body = json.dumps(r)
    dump_s = time.perf_counter() - t1
    arr = op.unpack_result(r) if name != "json lists" else None
    labels[name] = r["labels"] if arr is None else arr["labels"].tolist()
    print(json.dumps({"what": f"map_classify_tpu op, {n} items, one call, result as {name}", "devices": len(devices),
                      "op_wall_s": round(wall, 3), "items_per_s_wall": round(n / wall),
                      "device_ms": round(r["metrics"]["gpu_ms_total"], 1),
                      "items_per_s_device": round(n / r["metrics"]["gpu_ms_total"] * 1e3),
                      "result_json_bytes": len(body), "json_dumps_s": round(dump_s, 3),
                      "histogram_sum": sum(r["label_histogram"])}), flush=True)
assert labels["json lists"] == labels["packed"] == labels["result_uri"]
print(json.dumps({"what": "the three result forms carry identical labels", "ok": True}))
