#!/usr/bin/env python3
"""Generate tests/golden/reference_golden.json by IMPORTING the reference (read-only) and recording
what its own functions return for a fixed set of payloads.

Run in the build container only (the GPU box has no /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_reference_golden.py

The reference has no tests or fixtures of its own; these known-answer vectors are
its observed behaviour and pin the parts of the hot path that the reference does implement:
registry, envelope, top-k, chunking, task extraction, label parsing, worker-profile schema.
"""
import io
import json
import os
import sys
from contextlib import redirect_stdout

REF = "/root/reference"
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
for k in ("TASKS", "TPU_MODEL_PATH", "CUDA_MODEL_PATH"):
    os.environ.pop(k, None)

import numpy as np  # noqa: E402
import ops  # noqa: E402
from ops import map_classify_tpu as ref_classify  # noqa: E402
from ops import map_tokenize as ref_tok  # noqa: E402
from ops import _tpu_runtime as ref_rt  # noqa: E402
import ops_loader  # noqa: E402
import worker_sizing  # noqa: E402

out = {}

# ---- registry ------------------------------------------------------------------------------
reg = {}
reg["get_op_classify_is_run"] = ops.get_op("map_classify_tpu") is ref_classify.run
reg["load_ops_keys"] = sorted(ops_loader.load_ops(["map_classify_tpu", "map_tokenize", "echo"]).keys())
for name, label in (("nonexistent_op", "unknown"),):
    try:
        ops.get_op(name)
        reg[label] = None
    except ValueError as e:
        reg[label] = str(e).split(". ")[0]
os.environ["TASKS"] = "echo"
try:
    ops.get_op("map_classify_tpu")
    reg["disabled"] = None
except ValueError as e:
    reg["disabled"] = str(e)
reg["list_ops_tasks_echo"] = ops.list_ops()
os.environ["TASKS"] = "none"
reg["list_ops_tasks_none"] = ops.list_ops()
os.environ["TASKS"] = "ALL"
reg["list_ops_tasks_all_len_matches_map"] = len(ops.list_ops()) == len(ops.OP_TO_MODULE)
os.environ.pop("TASKS")
out["registry"] = reg

# ---- model path precedence -------------------------------------------------------------------
mp = {}
mp["default"] = ref_rt.get_model_path(None)
os.environ["TPU_MODEL_PATH"] = "/env/model.tflite"
mp["env"] = ref_rt.get_model_path(None)
mp["requested_over_env"] = ref_rt.get_model_path("/req/model.tflite")
mp["empty_string_falls_through"] = ref_rt.get_model_path("")
os.environ.pop("TPU_MODEL_PATH")
out["model_path"] = mp

# ---- _topk -----------------------------------------------------------------------------------
tk = []
rng = np.random.default_rng(7)
cases = [([0.1, 0.9, 0.5, 0.7], 3), ([0.1, 0.9], 5), ([0.3, 0.2, 0.1], 0), ([0.5], 1), ([1.0, -2.0, 3.5, 0.25, 3.25], -4)]
for _ in range(6):
    n = int(rng.integers(2, 40))
    cases.append((rng.permutation(n).astype(np.float32).tolist(), int(rng.integers(1, n + 3))))
for scores, k in cases:
    tk.append({"scores": scores, "k": k, "out": ref_classify._topk(np.array(scores, dtype=np.float32), k)})
out["topk"] = tk

# ---- classify envelope without an accelerator ---------------------------------------------------
env = []
for payload in ({"input": [0] * 10}, {"input": [0] * 10, "topk": 3, "model_path": "/x/y.tflite"}, {}, None,
                {"input": [1, 2, 3], "fallback_reason": "ignored"}):
    r = ref_classify.run(payload)
    r["elapsed_ms"] = "float" if isinstance(r.get("elapsed_ms"), float) else repr(type(r.get("elapsed_ms")))
    env.append({"payload": payload, "result": r, "keys": list(r.keys())})
try:
    ref_classify.run({"input": [0] * 10, "allow_fallback": False})
    raised = None
except Exception as e:
    raised = type(e).__name__
out["classify_envelope"] = {"cases": env, "allow_fallback_false_raises": raised}

# ---- map_tokenize ------------------------------------------------------------------------------
tc = []
payloads = [
    {"text": "abcdefghij", "chunk_size": 4},
    {"items": ["abcde", None, 123, ""], "chunk_size": 2},
    {"chunk_size": 0}, {"chunk_size": -1}, {"chunk_size": "4"}, {"chunk_size": 2.0},
    None, {}, {"text": ""}, {"data": "from data key", "chunk_size": 5},
    {"text": "", "data": "text empty so data wins", "chunk_size": 7},
    {"text": 5}, {"items": "not a list"}, {"items": None, "text": "items None falls to text", "chunk_size": 6},
    {"items": [], "chunk_size": 3},
    {"text": "ééé", "chunk_size": 2},
    {"text": "naïve café 中文字 😀😀 done", "chunk_size": 3},
    {"items": ["αβγδε", "x" * 2500, True, 1.5, ["nested"]]},
    {"items": ["line one\nline two", "\ttabbed"], "chunk_size": 1024},
    {"text": "x" * 1024}, {"text": "x" * 1025},
]
for p in payloads:
    tc.append({"payload": p, "result": ref_tok.map_tokenize_op(p)})
out["map_tokenize"] = tc

# ---- app.py helpers (the caller of the hot path) -----------------------------------------------
buf = io.StringIO()
with redirect_stdout(buf):
    import app  # noqa: E402
ap = {}
ap["extract"] = [list(app._extract_task({"id": "j1", "op": "echo", "payload": None, "job_epoch": 3})),
                 list(app._extract_task({"job_id": "j2", "op": "map_classify_tpu", "payload": {"input": [1]}}))]
ap["labels"] = app._parse_labels("zone=a, tpu ,x=1=2")
ap["ops_keys"] = sorted(app.OPS.keys())
ap["worker_profile_keys"] = sorted(app.WORKER_PROFILE.keys())
out["app"] = ap

# ---- worker_sizing schema ---------------------------------------------------------------------
ws = {}
os.environ["NVIDIA_VISIBLE_DEVICES"] = "none"
prof = worker_sizing.build_worker_profile()
ws["top_keys"] = sorted(prof.keys())
ws["cpu_keys"] = sorted(prof["cpu"].keys())
ws["gpu_hidden"] = prof["gpu"]
ws["tpu_keys"] = sorted(prof["tpu"].keys())
ws["workers_keys"] = sorted(prof["workers"].keys())
os.environ.pop("NVIDIA_VISIBLE_DEVICES")
# the CPU sizing rule, evaluated for fixed core counts through the env knobs it reads
rule = []
import unittest.mock as um
for cores in (1, 2, 4, 8, 16, 64, 256):
    with um.patch.object(worker_sizing, "psutil", None), um.patch("os.cpu_count", return_value=cores):
        c = worker_sizing._detect_cpu()
    rule.append({"cores": cores, "cpu": c})
ws["cpu_rule_no_psutil"] = rule
# nvidia-smi parser on canned output
with um.patch("subprocess.check_output", return_value="NVIDIA B200, 183359\nNVIDIA B200, 183359\nbad line\nX, notanumber\n"):
    ws["parse_smi"] = worker_sizing._parse_nvidia_smi()
    ws["detect_gpu"] = worker_sizing._detect_gpu()
out["worker_sizing"] = ws

dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_golden.json")
with open(dst, "w", encoding="utf-8") as f:
    json.dump(out, f, indent=1, ensure_ascii=False, sort_keys=True)
print("wrote", dst, os.path.getsize(dst), "bytes")
