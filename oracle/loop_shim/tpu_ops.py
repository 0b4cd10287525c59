"""`tpu_ops` as the agent loop imports it (app.py:118-123), backed by the CPU oracle.  TEST INFRASTRUCTURE.

Only bench.py's `--impl reference --workload agent_loop` leg puts this directory on an agent's
PYTHONPATH: it is BASELINE.md's B2 — the CPU restatement of the classifier (tokenizers + fp32 torch
BERT-base, oracle/bert_oracle.py; NOT the reference, whose own CPU path is a stub) driven through the
same lease / result loop as the GPU op, so that both arms pay the same protocol costs. The product's
`tpu_ops.py` is agent-tpu_b200/tpu_ops.py and never imports this.
"""
import os
import sys
import time

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (os.path.join(_ROOT, "oracle"), os.path.join(_ROOT, "agent-tpu_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

_STATE = {}


def map_classify_tpu(payload: dict) -> dict:
    import numpy as np
    import torch
    import bert_oracle as bo
    t0 = time.time()
    payload = payload or {}
    model, vocab = os.environ["CUDA_MODEL_PATH"], os.environ["CUDA_VOCAB_PATH"]
    if "model" not in _STATE:
        torch.set_num_threads(os.cpu_count() or 1)
        _STATE["model"] = bo.load_model(model)
    cfg, W = _STATE["model"]
    items = ["" if x is None else str(x) for x in payload.get("items", [])]
    k = max(1, min(int(payload.get("topk", 5)), cfg.num_labels))
    ids, mask = bo.tokenize(vocab, items, int(payload.get("max_len", 128)))
    logits = bo.forward(cfg, W, ids, mask, batch=int(payload.get("batch", 32)))
    probs = bo.softmax(logits.astype(np.float64)).astype(np.float32)
    labels = logits.argmax(-1)
    return {"op": "map_classify_tpu", "model_path": model, "topk": bo.topk(probs[0], k) if len(items) else [],
            "elapsed_ms": (time.time() - t0) * 1e3, "items_count": len(items), "labels": [int(v) for v in labels],
            "label_histogram": [int(v) for v in np.bincount(labels, minlength=cfg.num_labels)]}
