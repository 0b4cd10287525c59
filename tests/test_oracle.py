"""The oracle against its pins: the reference's golden top-k, the committed tokenizer / logits
vectors, and transformers' BertForSequenceClassification. CPU only."""
import json
import os

import numpy as np
import pytest
import torch

import bert_oracle as bo
import model_file as mf


def test_topk_restatement_matches_reference_golden(ref_golden):
    for case in ref_golden["topk"]:
        scores = np.array(case["scores"], dtype=np.float32)
        got = bo.topk(scores, case["k"])
        want = case["out"]
        assert len(got) == len(want)
        # order among exactly equal scores is unspecified in the reference (unstable sort)
        assert [g["score"] for g in got] == [w["score"] for w in want]
        if len(set(case["scores"])) == len(case["scores"]):
            assert [g["index"] for g in got] == [w["index"] for w in want]


def test_vocab_is_the_one_the_golden_vectors_were_made_with(golden_dir, vocab_sha):
    with open(os.path.join(golden_dir, "tokenizer_golden.json"), encoding="utf-8") as f:
        g = json.load(f)
    assert g["vocab_sha256"] == vocab_sha


def test_tokenizer_oracle_matches_committed_golden(golden_dir, vocab_path):
    with open(os.path.join(golden_dir, "tokenizer_golden.json"), encoding="utf-8") as f:
        g = json.load(f)
    for case in g["cases"]:
        ids, mask = bo.tokenize(vocab_path, g["texts"], case["max_len"])
        for row, m, want in zip(ids, mask, case["ids"]):
            n = int(m.sum())
            assert row[:n].tolist() == want
            assert not row[n:].any() and m[:n].all() and not m[n:].any()


def test_forward_matches_transformers_bert(tiny_model, vocab_path, sentences):
    cfg, W = bo.load_model(tiny_model)
    ids, mask = bo.tokenize(vocab_path, sentences[:24] + ["", "x"], 128)
    mine = bo.forward(cfg, W, ids, mask)
    model = bo.hf_model(cfg, W)
    with torch.no_grad():
        ref = model(input_ids=torch.from_numpy(ids).long(), attention_mask=torch.from_numpy(mask).long()).logits.numpy()
    assert np.abs(mine - ref).max() < 2e-5


def test_forward_matches_committed_golden_tiny(golden_dir, tiny_model, vocab_path):
    with open(os.path.join(golden_dir, "classify_tiny_texts.json"), encoding="utf-8") as f:
        meta = json.load(f)
    z = np.load(os.path.join(golden_dir, "classify_tiny_golden.npz"))
    cfg, W = bo.load_model(tiny_model)
    assert mf.config_dict(cfg) == meta["model"]
    ids, mask = bo.tokenize(vocab_path, meta["texts"], meta["max_len"])
    assert (mask.sum(1) == z["lens"]).all() and int(ids.astype(np.int64).sum()) == int(z["ids_crc"][0])
    assert np.abs(bo.forward(cfg, W, ids, mask) - z["logits"]).max() < 1e-4


def test_forward_matches_committed_golden_base_sample(golden_dir, base_model, vocab_path):
    with open(os.path.join(golden_dir, "classify_base_texts.json"), encoding="utf-8") as f:
        meta = json.load(f)
    z = np.load(os.path.join(golden_dir, "classify_base_golden.npz"))
    cfg, W = bo.load_model(base_model)
    pick = [0, 3, 9, 100, 255]
    ids, mask = bo.tokenize(vocab_path, [meta["texts"][i] for i in pick], meta["max_len"])
    assert np.abs(bo.forward(cfg, W, ids, mask) - z["logits"][pick]).max() < 1e-4


def test_bf16_emulation_predicts_the_stated_tolerance(base_model, vocab_path, sentences):
    """The 2e-2 logit tolerance (BASELINE.json) has to be reachable with bf16 activations at all."""
    cfg, W = bo.load_model(base_model)
    ids, mask = bo.tokenize(vocab_path, sentences[:8], 128)
    exact = bo.forward(cfg, W, ids, mask)
    emu = bo.forward(cfg, W, ids, mask, emulate_bf16=True)
    assert np.abs(exact - emu).max() < 2e-2


def test_chunk_oracle_reproduces_the_reference_map_tokenize(ref_golden):
    """oracle/chunk_oracle.py (bench.py's B3 baseline) against the outputs recorded from the
    reference's own ops/map_tokenize.py: identical dictionaries for all recorded payloads."""
    import chunk_oracle
    assert len(ref_golden["map_tokenize"]) >= 21
    for case in ref_golden["map_tokenize"]:
        assert chunk_oracle.map_tokenize(case["payload"]) == case["result"], case["payload"]
    assert chunk_oracle.map_tokenize(None) == {"ok": True, "tokens": [], "count": 0, "total_chars": 0}


def test_bart_encoder_maps_onto_the_model_file(tmp_path):
    """The encoder of `map_summarize`'s model (BART: ops/map_summarize.py:29-30) is this library's layer
    with different embeddings: a random-init transformers BartEncoder written as a .b2m file
    (tests/bart_encoder.py) and run through the fp32 oracle reproduces BartEncoder's own hidden
    states. The GPU counterpart is tests/test_gpu_parity_layers.py::test_bart_encoder_on_these_kernels."""
    import bart_encoder as be
    import bert_oracle as bo
    model = be.random_bart(layers=2)
    path, cfg = be.encoder_as_model_file(model, str(tmp_path / "bart_enc.b2m"))
    rng = np.random.default_rng(0)
    lens, S = [1, 5, 64, 128], 128
    ids = np.ones((len(lens), S), np.int32)
    mask = np.zeros((len(lens), S), np.uint8)
    for i, L in enumerate(lens):
        ids[i, :L] = rng.integers(3, cfg.vocab_size, L)
        mask[i, :L] = 1
    c, W = bo.load_model(path)
    _, hs = bo.forward(c, W, ids, mask, return_hidden=True)
    ref = be.bart_hidden_states(model, ids, mask)
    assert hs.shape == ref.shape == (3, 4, S, 768)
    assert np.abs(hs - ref).max() < 2e-5
