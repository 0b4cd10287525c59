#!/usr/bin/env python3
"""GPU box: map_tokenize config 2 — WordPiece tokenize-and-pad of 1M synthetic lines, inputs resident
in HBM, timed with CUDA events on the library's stream (b2_tokenize_device metrics)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
import torch
import synth
from ops import _cuda_runtime as rt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
tmp = "/tmp/b2tok"; os.makedirs(tmp, exist_ok=True)
vocab = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
pool = synth.SentenceGenerator().sentences(8192)
texts = (pool * (n // len(pool) + 1))[:n]
blob, off = rt.pack_items(texts)
eng = rt.Engine(None, vocab, max_batch=1, max_len=128)
d_blob = torch.from_numpy(blob.copy()).cuda(); d_off = torch.from_numpy(off).cuda()
d_ids = torch.empty((n, 128), dtype=torch.int32, device="cuda")
d_mask = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
for it in range(4):
    m = eng.tokenize_device(d_blob.data_ptr(), d_off.data_ptr(), n, blob.size, 128, d_ids.data_ptr(), d_mask.data_ptr())
    ms = m["gpu_ms_total"]
    alg = blob.size + 8 * n + 5 * 128 * n
    print(f"run {it}: {n} lines, {blob.size/1e6:.1f} MB text, {ms:.2f} ms -> {n/ms*1e3/1e6:.2f} M lines/s, "
          f"{alg/ms/1e6:.1f} GB/s algorithmic ({alg/1e6:.0f} MB)")
t = time.time(); ids, mask, _ = eng.tokenize(texts[:200000], 128); print("host API 200k lines wall", time.time() - t)
