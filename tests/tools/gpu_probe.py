#!/usr/bin/env python3
"""Developer probe (GPU box): tokenizer vs tokenizers, classifier vs the fp32 oracle, a timing."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import synth, model_file as mf
from ops import _cuda_runtime as rt
import bert_oracle as bo

tmp = "/tmp/b2probe"; os.makedirs(tmp, exist_ok=True)
vocab = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
tiny = mf.write_model_file(os.path.join(tmp, "tiny.b2m"), mf.BERT_TINY)
base = mf.write_model_file(os.path.join(tmp, "base.b2m"), mf.BERT_BASE)
gen = synth.SentenceGenerator()
texts = gen.sentences(2048)
edge = ["", " ", "Hello, World!", "Héllo Wörld — naïve café", "中文字 and 日本語 mixed", "a​b\x00c�d",
        "x" * 150 + " tail", "tab\tnew\nline\r\nend", "[CLS] literal [SEP] and [MASK] [PAD] [UNK] [cls]",
        "İstanbul ǅ ﬃ 한국어 😀 emoji", "##weird ## tokens #", "a" * 100, "b" * 101, "word " * 300,
        "ʒʒʒ unknown ћ letters", "3.14159 + 2,718 = ?", "　ideographic spaces nbsp"]
eng = rt.Engine(tiny, vocab, max_batch=256, max_len=512)
for ml in (128, 16, 512):
    tt = edge + texts[:500]
    ids, mask, _ = eng.tokenize(tt, ml)
    rids, rmask = bo.tokenize(vocab, tt, ml)
    bad = np.nonzero((ids != rids).any(1) | (mask != rmask).any(1))[0]
    print(f"tokenize max_len={ml}: {len(tt)} lines, mismatching lines: {len(bad)}")
    for i in bad[:5]:
        print("  line", i, repr(tt[i][:60])); print("   got ", ids[i][:24].tolist()); print("   want", rids[i][:24].tolist())
# classify tiny
out = eng.classify(texts[:300], 128, 3, want_logits=True)
cfg, W = bo.load_model(tiny)
rids, rmask = bo.tokenize(vocab, texts[:300], 128)
ref = bo.forward(cfg, W, rids, rmask)
print("tiny logits max abs diff", np.abs(out["logits"] - ref).max(), "labels equal", (out["labels"] == ref.argmax(1)).mean(), out["metrics"])
eng.close()
eng = rt.Engine(base, vocab, max_batch=1024, max_len=128)
out = eng.classify(texts[:64], 128, 5, want_logits=True)
cfg, W = bo.load_model(base)
rids, rmask = bo.tokenize(vocab, texts[:64], 128)
t = time.time(); ref = bo.forward(cfg, W, rids, rmask); cpu_s = time.time() - t
print("base logits max abs diff", np.abs(out["logits"] - ref).max(), "labels equal", (out["labels"] == ref.argmax(1)).mean(), "hist", out["histogram"].tolist(), "cpu s/item", cpu_s / 64)
eng.set_profiling(True)
big = texts * 2
for _ in range(2):
    t = time.time(); out = eng.classify(big, 128, 5); dt = time.time() - t
    m = out["metrics"]
    print(f"base {len(big)} items: wall {dt*1e3:.1f} ms, gpu {m['gpu_ms_total']:.1f} ms -> {len(big)/m['gpu_ms_total']*1e3:.0f} items/s (gpu), launches {m['kernel_launches']}")
for k, v in m.get("kernels", {}).items():
    print(f"   {k:14s} {v['ms']:9.3f} ms  {v['launches']:5d} launches")
