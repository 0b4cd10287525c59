#!/usr/bin/env python3
"""GPU box: BASELINE config 4 shape — long documents, max_len 512 (attention-heavy), one GPU."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import synth, model_file as mf
from ops import _cuda_runtime as rt

tmp = "/tmp/b2s512"; os.makedirs(tmp, exist_ok=True)
vocab = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
base = mf.write_model_file(os.path.join(tmp, "base.b2m"), mf.BERT_BASE)
gen = synth.SentenceGenerator()
docs = gen.sentences(512, min_tokens=256, max_tokens=510, seed=99)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
texts = (docs * (n // len(docs) + 1))[:n]
eng = rt.Engine(base, vocab, max_batch=256, max_len=512)
eng.set_profiling(True)
for it in range(3):
    t = time.time(); out = eng.classify(texts, 512, 5); dt = time.time() - t
    m = out["metrics"]
    print(f"S=512 {n} docs: wall {dt*1e3:.0f} ms, gpu {m['gpu_ms_total']:.0f} ms -> {n/m['gpu_ms_total']*1e3:.0f} items/s, "
          f"tokens/item {m['tokens']/n:.0f}")
print({k: round(v["ms"], 1) for k, v in m["kernels"].items()})
if len(sys.argv) > 2:
    import bert_oracle as bo
    cfg, W = bo.load_model(base)
    ids, mask = bo.tokenize(vocab, texts[:8], 512)
    ref = bo.forward(cfg, W, ids, mask, batch=4)
    o = eng.classify(texts[:8], 512, 5, want_logits=True)
    print("S=512 base logits max abs diff vs oracle", np.abs(o["logits"] - ref).max())
