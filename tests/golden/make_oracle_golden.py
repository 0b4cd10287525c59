#!/usr/bin/env python3
"""Generate the oracle's golden vectors (committed) for the parts the reference does not implement:
token ids from tokenizers.BertWordPieceTokenizer and fp32 logits from oracle/bert_oracle.forward.

    python tests/golden/make_oracle_golden.py

Inputs are pure functions of seeds (agent-tpu_b200/synth.py, model_file.py); the texts themselves
are stored next to the outputs so the fixtures do not depend on generator stability.
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "agent-tpu_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bert_oracle as bo  # noqa: E402
import model_file as mf  # noqa: E402
import synth  # noqa: E402

EDGE_TEXTS = [
    "", " ", "   \t\n  ", "Hello, World!", "Héllo Wörld — naïve café", "中文字 and 日本語 mixed 汉字",
    "a​b\x00c�d", "x" * 150 + " tail", "tab\tnew\nline\r\nend",
    "[CLS] literal [SEP] and [MASK] [PAD] [UNK] [cls] x[SEP]y [[SEP]] [SEP", "İstanbul ǅ ﬃ 한국어 😀 emoji",
    "##weird ## tokens #", "a" * 100, "b" * 101, "c" * 99 + "é", "word " * 300,
    "ʒʒʒ unknown ћ letters", "3.14159 + 2,718 = ?", "　ideographic spaces nbsp",
    "ＦＵＬＬwidth ｔｅｘｔ", "Ǆǅǆ ǈ ß ẞ ſ", "é ǟ ộ combining", "́leading mark",
    "ΑΒΓ αβγ ΣΊΣΥΦΟΣ", "Ünïcödé ÀÉÎÕÜ ÇŁØ", "end.", "...", "don't stop-me_now", "a" * 400,
    "pad" + " " * 200 + "gap", "tiotstiol pleslaitcean zzzqqq", "\U0001F600\U0001F601", "𠀀 ext-B ideograph",
    "ก็ Thai กำ marks", "क्षत्रिय देवनागरी", "\x7f\x80\x9f c1 controls", " line sep", "﻿bom",
]


def main():
    tmp = tempfile.mkdtemp(prefix="b2golden")
    vocab_path = synth.write_vocab(os.path.join(tmp, "vocab.txt"))
    vocab_sha = hashlib.sha256(open(vocab_path, "rb").read()).hexdigest()
    gen = synth.SentenceGenerator()
    sentences = gen.sentences(256)
    long_docs = gen.sentences(16, min_tokens=256, max_tokens=510, seed=99)

    # ---- tokenizer golden: ids without padding, for several max_len -------------------------------
    tok = {"vocab_sha256": vocab_sha, "cases": []}
    texts = EDGE_TEXTS + sentences[:96] + long_docs[:4]
    for max_len in (128, 16, 512, 2):
        ids, mask = bo.tokenize(vocab_path, texts, max_len)
        tok["cases"].append({"max_len": max_len, "ids": [row[: int(m.sum())].tolist() for row, m in zip(ids, mask)]})
    tok["texts"] = texts
    with open(os.path.join(HERE, "tokenizer_golden.json"), "w", encoding="utf-8") as f:
        json.dump(tok, f, ensure_ascii=False)

    # ---- classifier golden: fp32 oracle logits ---------------------------------------------------
    for name, cfg, n in (("tiny", mf.BERT_TINY, 256), ("base", mf.BERT_BASE, 256)):
        path = mf.write_model_file(os.path.join(tmp, name + ".b2m"), cfg)
        c, W = bo.load_model(path)
        t = EDGE_TEXTS[:8] + sentences[: n - 8]
        ids, mask = bo.tokenize(vocab_path, t, 128)
        logits = bo.forward(c, W, ids, mask)
        np.savez_compressed(os.path.join(HERE, f"classify_{name}_golden.npz"), logits=logits,
                            lens=mask.sum(1).astype(np.int32), ids_crc=np.array([int(ids.astype(np.int64).sum())]))
        with open(os.path.join(HERE, f"classify_{name}_texts.json"), "w", encoding="utf-8") as f:
            json.dump({"texts": t, "model": mf.config_dict(c), "vocab_sha256": vocab_sha, "max_len": 128}, f,
                      ensure_ascii=False)
        print(name, logits.shape, "labels", np.bincount(logits.argmax(1), minlength=c.num_labels).tolist())
    # ---- wide-logit variant of the base model -----------------------------------------------------
    # Random-init BERT-base logits barely move between inputs (per-class std 0.02-0.05 over these
    # texts), so an absolute 2e-2 tolerance and an arg-max that is nearly constant say little. This
    # variant keeps the encoder and pooler and replaces the head: Wc' = scale * Wc and
    # bc' = -Wc' . mean(pooled), which centres every class over the text set, so that logits span
    # several units and every class wins somewhere. Only scale and bc' are stored (the test rebuilds
    # the kind-0 model file from them); tolerances scale with `scale`.
    path = os.path.join(tmp, "base.b2m")
    c, W = bo.load_model(path)
    t = EDGE_TEXTS[:8] + sentences[: 256 - 8]
    ids, mask = bo.tokenize(vocab_path, t, 128)
    _, hs = bo.forward(c, W, ids, mask, return_hidden=True)
    import torch
    x0 = torch.from_numpy(hs[-1][:, 0, :])
    pooled = torch.tanh(x0 @ W["Wp"].T + W["bp"])
    centred = (pooled - pooled.mean(0, keepdim=True)) @ W["Wc"].T
    scale = float(np.float32(4.0 / float(centred.abs().max())))       # largest |logit| becomes ~4
    Wc2 = (W["Wc"] * np.float32(scale)).contiguous()
    bc2 = -(Wc2 @ pooled.mean(0))
    logits = (pooled @ Wc2.T + bc2).numpy().astype(np.float32)
    wins = np.bincount(logits.argmax(1), minlength=c.num_labels)
    print("wide: scale %.2f, logits in [%.2f, %.2f], per-class std %.2f..%.2f, wins %s" % (
        scale, logits.min(), logits.max(), logits.std(0).min(), logits.std(0).max(), wins.tolist()))
    assert (wins > 0).all() and logits.max() >= 3 and logits.min() <= -3
    np.savez_compressed(os.path.join(HERE, "classify_wide_golden.npz"), logits=logits, scale=np.float32(scale),
                        bc=bc2.numpy().astype(np.float32), lens=mask.sum(1).astype(np.int32))
    # S = 512 (long documents) on the tiny model
    path = os.path.join(tmp, "tiny.b2m")
    c, W = bo.load_model(path)
    ids, mask = bo.tokenize(vocab_path, long_docs, 512)
    logits = bo.forward(c, W, ids, mask, batch=8)
    np.savez_compressed(os.path.join(HERE, "classify_tiny512_golden.npz"), logits=logits, lens=mask.sum(1).astype(np.int32))
    with open(os.path.join(HERE, "classify_tiny512_texts.json"), "w", encoding="utf-8") as f:
        json.dump({"texts": long_docs, "model": mf.config_dict(c), "vocab_sha256": vocab_sha, "max_len": 512}, f,
                  ensure_ascii=False)
    print("vocab sha", vocab_sha)


if __name__ == "__main__":
    main()
