"""map_classify parity on the GPU, through the C ABI and through the op.

Tolerances are the ones BASELINE.json states: token ids / masks bit-exact (test_gpu_tokenizer.py);
logits within 2e-2 max-abs of the fp32 oracle; predicted labels identical wherever the oracle's
top-2 margin exceeds 0.05. The oracle is oracle/bert_oracle.py (pinned to transformers'
BertForSequenceClassification; NOT the reference, which has no model — SURVEY.md §0)."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2
MARGIN = 0.05


def _load_golden(golden_dir, name):
    with open(os.path.join(golden_dir, f"classify_{name}_texts.json"), encoding="utf-8") as f:
        meta = json.load(f)
    z = np.load(os.path.join(golden_dir, f"classify_{name}_golden.npz"))
    return meta, z


def _check_against(logits_ref, out, k):
    logits = out["logits"]
    assert np.abs(logits - logits_ref).max() <= LOGIT_TOL
    srt = np.sort(logits_ref, axis=1)
    sure = (srt[:, -1] - srt[:, -2]) > MARGIN
    assert (out["labels"][sure] == logits_ref.argmax(1)[sure]).all()
    # head consistency: labels / top-k / scores / histogram all derive from the device logits
    assert (out["labels"] == logits.argmax(1)).all()
    assert (out["top_idx"][:, 0] == out["labels"]).all()
    e = np.exp(logits - logits.max(1, keepdims=True))
    probs = e / e.sum(1, keepdims=True)
    assert np.abs(np.take_along_axis(probs, out["top_idx"], 1) - out["top_score"]).max() < 1e-5
    assert (np.diff(out["top_score"], axis=1) <= 1e-7).all()
    assert (out["histogram"] == np.bincount(out["labels"], minlength=logits.shape[1])).all()
    assert out["top_idx"].shape[1] == k


@pytest.fixture(scope="module")
def tiny_engine(tiny_model, vocab_path):
    from ops import _cuda_runtime as rt
    eng = rt.Engine(tiny_model, vocab_path, max_batch=64, max_len=512)
    yield eng
    eng.close()


@pytest.fixture(scope="module")
def base_engine(base_model, vocab_path):
    from ops import _cuda_runtime as rt
    eng = rt.Engine(base_model, vocab_path, max_batch=1024, max_len=128)
    yield eng
    eng.close()


def test_tiny_model_against_committed_golden(tiny_engine, golden_dir):
    meta, z = _load_golden(golden_dir, "tiny")
    out = tiny_engine.classify(meta["texts"], 128, 3, want_logits=True)   # 256 items = 4 batches of 64
    _check_against(z["logits"], out, 3)
    assert out["metrics"]["tokens"] == int(z["lens"].sum()) and out["metrics"]["batches"] == 4


def test_tiny_model_long_documents_s512(tiny_engine, golden_dir):
    meta, z = _load_golden(golden_dir, "tiny512")
    out = tiny_engine.classify(meta["texts"], 512, 5, want_logits=True)
    _check_against(z["logits"], out, 5)


def test_base_model_against_committed_golden(base_engine, golden_dir):
    meta, z = _load_golden(golden_dir, "base")
    out = base_engine.classify(meta["texts"], 128, 5, want_logits=True)
    _check_against(z["logits"], out, 5)
    assert len(set(out["labels"].tolist())) >= 2      # the fixture exercises more than one class


def test_base_model_against_live_oracle(base_engine, base_model, vocab_path, sentences):
    import bert_oracle as bo
    texts = sentences[1000:1048]
    cfg, W = bo.load_model(base_model)
    ids, mask = bo.tokenize(vocab_path, texts, 128)
    ref = bo.forward(cfg, W, ids, mask)
    out = base_engine.classify(texts, 128, 5, want_logits=True)
    _check_against(ref, out, 5)


def test_explicit_tensor_model_file_equals_synthetic(tiny_model, vocab_path, work_dir, sentences):
    """kind 0 (f32 tensors in the file) loads to the same network as kind 1 (seed): this also
    proves the device-side synthetic initialiser is bit-identical to the host twin."""
    import model_file as mf
    from ops import _cuda_runtime as rt
    cfg = mf.read_header(tiny_model)
    p = mf.write_model_file(os.path.join(work_dir, "tiny_explicit.b2m"), cfg, tensors=mf.load_tensors(tiny_model))
    a = rt.Engine(tiny_model, vocab_path, max_batch=32, max_len=128)
    b = rt.Engine(p, vocab_path, max_batch=32, max_len=128)
    try:
        oa = a.classify(sentences[:40], 128, 2, want_logits=True)
        ob = b.classify(sentences[:40], 128, 2, want_logits=True)
        assert (oa["logits"] == ob["logits"]).all() and (oa["top_idx"] == ob["top_idx"]).all()
    finally:
        a.close(); b.close()


def test_hidden_size_that_is_not_a_multiple_of_256(vocab_path, work_dir, sentences):
    """hidden 384 (6 heads), FFN 1152: the GEMMs' last 256-wide column tile is half empty, so some
    epilogue quarters sit out every other tile — their residual barrier must keep its own phase (it
    used to follow the tile counter). Logits and every layer against the fp32 oracle."""
    import bert_oracle as bo
    import model_file as mf
    from ops import _cuda_runtime as rt
    cfg = mf.ModelConfig(hidden=384, layers=3, heads=6, intermediate=1152, num_labels=7, init_std=0.05,
                         classifier_std=0.1, seed=99)
    path = mf.write_model_file(os.path.join(work_dir, "h384.b2m"), cfg)
    texts = sentences[:48]
    c, W = bo.load_model(path)
    ids, mask = bo.tokenize(vocab_path, texts, 128)
    ref, ref_hidden = bo.forward(c, W, ids, mask, return_hidden=True)
    eng = rt.Engine(path, vocab_path, max_batch=48, max_len=128)
    try:
        out = eng.classify(texts, 128, 3, want_logits=True)
        hidden, _ = eng.debug_hidden_states(ids, mask)
    finally:
        eng.close()
    live = mask.astype(bool)
    for l in range(cfg.layers):
        rel = np.abs(hidden[l][live] - ref_hidden[l][live]).max() / np.abs(ref_hidden[l][live]).max()
        assert rel <= 0.02, (l, rel)
    # this model's head is initialised ten times wider than BERT-base's (logits span +-1.6), so the
    # stated 2e-2 is taken relative to the span
    assert np.abs(out["logits"] - ref).max() <= LOGIT_TOL * max(1.0, float(np.abs(ref).max()))
    assert (out["labels"] == out["logits"].argmax(1)).all()


def test_batch_size_and_order_independence(tiny_model, vocab_path, sentences):
    from ops import _cuda_runtime as rt
    texts = sentences[:150] + [""] + sentences[150:200]
    a = rt.Engine(tiny_model, vocab_path, max_batch=64, max_len=128)
    b = rt.Engine(tiny_model, vocab_path, max_batch=7, max_len=128)
    try:
        oa = a.classify(texts, 128, 2, want_logits=True)
        ob = b.classify(texts, 128, 2, want_logits=True)
        assert (oa["logits"] == ob["logits"]).all() and (oa["histogram"] == ob["histogram"]).all()
        perm = np.random.default_rng(0).permutation(len(texts))
        oc = a.classify([texts[i] for i in perm], 128, 2, want_logits=True)
        assert (oc["logits"] == oa["logits"][perm]).all()
        assert (oc["histogram"] == oa["histogram"]).all()
    finally:
        a.close(); b.close()


def test_two_lanes_equal_one_lane_bit_for_bit(tiny_model, vocab_path, sentences, monkeypatch):
    """Batches alternating between the two buffer sets / streams (default) against B2_ONE_LANE=1, over
    enough batches that every lane is reused many times, text and ids entry points: identical bits."""
    from ops import _cuda_runtime as rt
    texts = sentences[:1500] + ["", "x"] + sentences[1500:2048]
    two = rt.Engine(tiny_model, vocab_path, max_batch=64, max_len=128)
    monkeypatch.setenv("B2_ONE_LANE", "1")
    one = rt.Engine(tiny_model, vocab_path, max_batch=64, max_len=128)
    monkeypatch.delenv("B2_ONE_LANE")
    try:
        assert two.info.activation_bytes > 1.9 * one.info.activation_bytes
        a = two.classify(texts, 128, 3, want_logits=True)
        b = one.classify(texts, 128, 3, want_logits=True)
        for key in ("logits", "top_idx", "top_score", "labels", "histogram"):
            assert (a[key] == b[key]).all(), key
        assert a["metrics"]["tokens"] == b["metrics"]["tokens"] and a["metrics"]["batches"] == 33
        ids, mask, _ = two.tokenize(texts, 128)
        c = two.classify_ids(ids, mask, 3, want_logits=True)
        assert (c["logits"] == a["logits"]).all() and (c["histogram"] == a["histogram"]).all()
        assert c["metrics"]["tokens"] == a["metrics"]["tokens"]
    finally:
        two.close(); one.close()


def test_million_item_job_properties(base_engine, sentences):
    """BASELINE config 3 at full size: 1,000,000 items in batches of 1,024 (977 batches over both lanes,
    31 tokenizer runs). The input repeats a 2,048-sentence block, so must every output."""
    n, period = 1_000_000, len(sentences)
    texts = (sentences * (n // period + 1))[:n]
    out = base_engine.classify(texts, 128, 5)
    m = out["metrics"]
    assert m["items"] == n and m["batches"] == 977
    assert int(out["histogram"].sum()) == n
    labels = out["labels"]
    assert (labels == np.tile(labels[:period], n // period + 1)[:n]).all()
    assert (out["top_idx"] == np.tile(out["top_idx"][:period], (n // period + 1, 1))[:n]).all()
    assert (out["top_score"] == np.tile(out["top_score"][:period], (n // period + 1, 1))[:n]).all()
    assert (out["histogram"] == np.bincount(labels, minlength=out["histogram"].size)).all()
    base = base_engine.classify(sentences, 128, 5)
    assert (base["labels"] == labels[:period]).all() and (base["top_score"] == out["top_score"][:period]).all()
    assert m["tokens"] == base["metrics"]["tokens"] * (n // period) + \
        base_engine.classify(texts[:n % period], 128, 5)["metrics"]["tokens"]


def test_long_document_job_properties(tiny_model, vocab_path):
    """BASELINE config 4's shape (S = 512, streaming attention kernel) over many batches and both
    lanes: 6,000 documents of 200-510 tokens repeating a 250-document block; outputs must repeat too,
    and match a one-lane engine bit for bit."""
    import synth
    from ops import _cuda_runtime as rt
    block = synth.SentenceGenerator().sentences(250, min_tokens=200, max_tokens=510)
    texts = block * 24
    eng = rt.Engine(tiny_model, vocab_path, max_batch=64, max_len=512)
    try:
        out = eng.classify(texts, 512, 3, want_logits=True)
        assert out["metrics"]["batches"] == 94 and int(out["histogram"].sum()) == len(texts)
        assert out["metrics"]["tokens"] > 200 * len(texts)
        assert (out["logits"] == np.tile(out["logits"][:250], (24, 1))).all()
        assert (out["labels"] == np.tile(out["labels"][:250], 24)).all()
        ref = eng.classify(block, 512, 3, want_logits=True)
        assert (ref["logits"] == out["logits"][:250]).all()
    finally:
        eng.close()


def test_packed_rows_equal_padded_rows_bit_for_bit(tiny_model, vocab_path, sentences, monkeypatch):
    """Padding-free token rows (default) against B2_FORCE_PADDED=1: same logits to the last bit."""
    from ops import _cuda_runtime as rt
    texts = sentences[:100] + ["", "x"] + sentences[100:140]
    a = rt.Engine(tiny_model, vocab_path, max_batch=48, max_len=512)
    monkeypatch.setenv("B2_FORCE_PADDED", "1")
    b = rt.Engine(tiny_model, vocab_path, max_batch=48, max_len=512)
    monkeypatch.delenv("B2_FORCE_PADDED")
    try:
        assert a.info.packed_rows == 1 and b.info.packed_rows == 0
        for ml in (128, 512):
            oa = a.classify(texts, ml, 3, want_logits=True)
            ob = b.classify(texts, ml, 3, want_logits=True)
            assert (oa["logits"] == ob["logits"]).all() and (oa["top_idx"] == ob["top_idx"]).all()
            assert (oa["histogram"] == ob["histogram"]).all()
    finally:
        a.close(); b.close()


def test_folded_layernorm_agrees_with_explicit_layernorm(base_model, vocab_path, sentences, monkeypatch):
    """LayerNorm folded into the GEMM epilogues (default) against B2_NO_LN_FOLD=1, which runs the
    LayerNorm kernels: two roundings of the same network, so close but not bit-identical, and both
    inside the stated tolerance of the fp32 oracle."""
    import bert_oracle as bo
    from ops import _cuda_runtime as rt
    texts = sentences[300:364]
    a = rt.Engine(base_model, vocab_path, max_batch=64, max_len=128)
    monkeypatch.setenv("B2_NO_LN_FOLD", "1")
    b = rt.Engine(base_model, vocab_path, max_batch=64, max_len=128)
    monkeypatch.delenv("B2_NO_LN_FOLD")
    try:
        assert a.info.ln_folded == 1 and b.info.ln_folded == 0
        oa = a.classify(texts, 128, 5, want_logits=True)
        ob = b.classify(texts, 128, 5, want_logits=True)
    finally:
        a.close(); b.close()
    cfg, W = bo.load_model(base_model)
    ids, mask = bo.tokenize(vocab_path, texts, 128)
    ref = bo.forward(cfg, W, ids, mask)
    ea, eb = np.abs(oa["logits"] - ref).max(), np.abs(ob["logits"] - ref).max()
    assert ea <= LOGIT_TOL and eb <= LOGIT_TOL, (ea, eb)
    assert np.abs(oa["logits"] - ob["logits"]).max() <= LOGIT_TOL


def test_ids_entry_point_equals_text_entry_point(tiny_engine, sentences):
    texts = sentences[:70]
    ids, mask, _ = tiny_engine.tokenize(texts, 128)
    a = tiny_engine.classify(texts, 128, 3, want_logits=True)
    b = tiny_engine.classify_ids(ids, mask, 3, want_logits=True)
    assert (a["logits"] == b["logits"]).all() and (a["top_idx"] == b["top_idx"]).all()


def test_limits_and_errors(tiny_engine):
    from ops import _cuda_runtime as rt
    with pytest.raises(rt.B2Error):
        tiny_engine.classify(["x"], 128, 99)            # topk > num_labels
    with pytest.raises(rt.B2Error):
        tiny_engine.classify(["x"], 1024, 1)            # beyond the handle's max_len
    with pytest.raises(rt.B2Error):
        tiny_engine.classify(["x"], 100, 1)             # not a multiple of 128
    out = tiny_engine.classify([], 128, 2)
    assert out["top_idx"].shape == (0, 2) and out["histogram"].sum() == 0


def test_full_batch_job_properties(base_engine, sentences):
    """Eight full 1,024-item batches (BASELINE config 3's batch shape): histogram sums to n, rows of a
    repeated input repeat, profiling accounts for every launch."""
    texts = (sentences * 4)[:8192]
    base_engine.set_profiling(True)
    try:
        out = base_engine.classify(texts, 128, 5)
    finally:
        base_engine.set_profiling(False)
    n = len(sentences)
    assert out["histogram"].sum() == len(texts)
    assert (out["labels"][:n] == out["labels"][n:2 * n]).all()
    assert (out["top_score"][:n] == out["top_score"][2 * n:3 * n]).all()
    m = out["metrics"]
    assert m["batches"] == 8 and m["items"] == 8192
    # timed segments: one tokenizer run for the job (up to 32,768 items per run), then per batch
    # embed, 11 layers x {QKV, attention, out-proj, row statistics, FFN1, FFN2}, the last layer's 7
    # ([CLS]-only tail with its two LayerNorms), pooler, head
    per_batch = 1 + 11 * 6 + 7 + 2
    assert sum(v["launches"] for v in m["kernels"].values()) == 1 + 8 * per_batch
    assert m["kernel_launches"] >= 1 + 8 * per_batch
    assert m["kernels"]["tokenize"]["launches"] == 1
    assert 0 < sum(v["ms"] for v in m["kernels"].values()) <= m["gpu_ms_total"] * 1.05


def test_op_envelope_on_gpu(base_model, vocab_path, clean_env, sentences, golden_dir):
    from ops import _cuda_runtime as rt
    from ops import map_classify_tpu as op
    rt.reset_cache()
    clean_env.setenv("CUDA_MODEL_PATH", base_model)
    texts = sentences[:33]
    r = op.run({"items": texts, "topk": 3, "return_results": True})
    assert list(r.keys())[:4] == ["op", "model_path", "topk", "elapsed_ms"]      # v0 keys first, same order
    assert r["op"] == "map_classify_tpu" and r["model_path"] == base_model
    assert len(r["topk"]) == 3 and all(set(t) == {"index", "score"} for t in r["topk"])
    assert all(type(t["index"]) is int and type(t["score"]) is float for t in r["topk"])
    assert r["items_count"] == 33 and len(r["labels"]) == 33 and sum(r["label_histogram"]) == 33
    assert r["results"][0]["topk"] == r["topk"]
    json.dumps(r)                                                               # pure JSON (app.py:146)
    # topk is clamped like the reference's _topk
    assert len(op.run({"items": texts[:2], "topk": 0})["topk"]) == 1
    assert len(op.run({"items": texts[:2], "topk": 999})["topk"]) == 16
    # the handle is cached by model path (ops/_tpu_runtime.py:42-43)
    h1 = rt.get_cuda_handle(base_model); h2 = rt.get_cuda_handle(base_model)
    assert h1.interpreter is h2.interpreter
    # v0 `input`: one pre-tokenised row
    ids, mask, _ = h1.interpreter.tokenize(texts[:1], 128)
    v0 = op.run({"input": ids[0].tolist(), "topk": 3})
    assert list(v0.keys()) == ["op", "model_path", "topk", "elapsed_ms"]
    assert [t["index"] for t in v0["topk"]] == [t["index"] for t in r["topk"]]
    # error routes (ops/map_classify_tpu.py:63-67, 84-90)
    bad = op.run({"input": [1, 2, 3]})
    assert bad["fallback"] == "cpu" and bad["topk"] == [] and bad["reason"] == \
        "Input size mismatch. Got 3, expected 128 for shape (1, 128)."
    assert op.run({})["reason"] == 'payload missing required key: "input"'
    with pytest.raises(ValueError):
        op.run({"input": [1, 2, 3], "allow_fallback": False})
    missing = op.run({"items": ["x"], "model_path": "/nonexistent/model.b2m"})
    assert missing["reason"] == "CUDA model not found: /nonexistent/model.b2m"
    rt.reset_cache()


def test_worker_profile_reports_this_gpu(clean_env, base_engine):
    import worker_sizing as ws
    from ops import _cuda_runtime as rt
    gpu = ws.build_worker_profile()["gpu"]
    assert gpu["gpu_present"] and gpu["gpu_count"] >= 1 and gpu["max_gpu_workers"] == gpu["gpu_count"]
    d = gpu["devices"][0]
    assert "B200" in d["name"] and d["sm_count"] == 148 and d["total_memory_bytes"] > 150 * 2 ** 30
    # capacity facts a controller can size `batch` x `max_len` with
    assert 0 < d["free_memory_bytes"] <= d["total_memory_bytes"]
    lim = d["classify_limits"]
    assert lim["max_batch_at_128"] == 4096 and lim["max_batch_at_512"] >= 1024      # > 100 GB free on this box
    # the per-row figure is the handle's own allocation: 1,024 x 128 rows of BERT-base
    rows = base_engine.info.max_batch * base_engine.info.max_len
    assert abs(base_engine.info.activation_bytes - rows * lim["bytes_per_token_row"]) < 0.02 * base_engine.info.activation_bytes
    if gpu["gpu_count"] > 1:
        link = gpu["interconnect"]
        assert link["gpu_count"] == gpu["gpu_count"] and link["all_to_all_peer_access"] and link["nvlink_all_to_all"]
        assert not any(link["peer_access"][i][i] for i in range(gpu["gpu_count"]))
    else:
        assert "interconnect" not in gpu
    free, total = rt.device_memory(0)
    assert total == d["total_memory_bytes"] or abs(total - d["total_memory_bytes"]) < 2 ** 30
