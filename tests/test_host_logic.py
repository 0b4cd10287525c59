"""Host-side pieces that need no GPU: synthetic data, model file, packing, sharding."""
import os

import numpy as np
import pytest

import model_file as mf
import synth
from ops import _cuda_runtime as rt
from ops import _shard


def test_vocab_layout():
    v = synth.build_vocab()
    assert len(v) == 30522 and len(set(v)) == 30522
    assert (v[0], v[100], v[101], v[102], v[103]) == ("[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]")


def test_sentences_are_deterministic_and_fit():
    g = synth.SentenceGenerator()
    a, b = g.sentences(50), g.sentences(50)
    assert a == b and all(s.endswith(".") for s in a)


def test_pack_items_layout():
    blob, off = rt.pack_items(["ab", "", "é", "\ud800"])
    assert off.tolist() == [0, 2, 2, 4, 7] and blob.dtype == np.uint8 and blob.size == 7
    blob, off = rt.pack_items([])
    assert off.tolist() == [0] and blob.size == 0
    assert synth.pack_items(["ab", "é"])[1].tolist() == [0, 2, 4]


def test_model_file_roundtrip_and_synthetic_stats(tmp_path):
    p = mf.write_model_file(str(tmp_path / "m.b2m"), mf.BERT_TINY)
    cfg = mf.read_header(p)
    assert (cfg.hidden, cfg.layers, cfg.heads, cfg.num_labels, cfg.kind) == (128, 2, 2, 5, 1)
    assert os.path.getsize(p) == mf.HEADER_SIZE
    t = mf.load_tensors(p)
    names = [n for n, *_ in mf.tensor_specs(cfg)]
    assert list(t.keys()) == names
    w = t["word_emb"]
    assert abs(float(w.std()) - cfg.init_std) < 2e-3 and abs(float(w.mean())) < 1e-3
    assert abs(float(t["emb_ln_g"].mean()) - 1.0) < 0.02
    # same seed -> same bits; another tensor id -> another stream
    a = mf.synthetic_tensor(1234, 0, 1000, np.float32(1e-6), np.float32(0))
    assert (a == mf.synthetic_tensor(1234, 0, 1000, np.float32(1e-6), np.float32(0))).all()
    assert (a != mf.synthetic_tensor(1234, 1, 1000, np.float32(1e-6), np.float32(0))).any()
    # explicit-tensor files carry the same values
    p2 = mf.write_model_file(str(tmp_path / "m2.b2m"), cfg, tensors=t)
    t2 = mf.load_tensors(p2)
    assert mf.read_header(p2).kind == 0 and all((t[k] == t2[k]).all() for k in t)


def test_slice_bounds_cover_without_overlap():
    for n in (0, 1, 7, 8, 1000003):
        for w in (1, 2, 4, 8):
            cuts = [_shard.slice_bounds(n, r, w) for r in range(w)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(w - 1))
            assert max(hi - lo for lo, hi in cuts) - min(hi - lo for lo, hi in cuts) <= 1


def test_shard_is_inert_without_process_group():
    assert _shard.my_slice(10) == (0, 10, 0, 1)
    h, m = _shard.reduce_job(np.arange(4), {"items": 3})
    assert h.tolist() == [0, 1, 2, 3] and m == {"items": 3}


def test_slice_bounds_partition_any_job():
    """Contiguous slices cover [0, n) exactly once for every job size and rank count, and differ in
    size by at most one item."""
    from hypothesis import given, settings, strategies as st
    from ops import _shard

    @settings(max_examples=200, deadline=None)
    @given(st.integers(0, 10_000_000), st.integers(1, 64))
    def check(n, world):
        edges = [_shard.slice_bounds(n, r, world) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
        sizes = [hi - lo for lo, hi in edges]
        assert min(sizes) >= 0 and max(sizes) - min(sizes) <= 1

    check()


def test_pack_items_layout_any_text():
    """pack_items: UTF-8 bytes back to back, offsets[i+1] - offsets[i] = byte length of item i, for
    ASCII-only jobs (one encode) and mixed ones alike."""
    from hypothesis import given, settings, strategies as st
    from ops._cuda_runtime import pack_items

    @settings(max_examples=200, deadline=None)
    @given(st.lists(st.text(max_size=40), max_size=30))
    def check(items):
        blob, off = pack_items(items)
        enc = [s.encode("utf-8", "surrogatepass") for s in items]
        assert blob.tobytes() == b"".join(enc)
        assert off.tolist() == [0] + list(__import__("itertools").accumulate(len(e) for e in enc))

    check()


def test_packed_and_file_results_round_trip(tmp_path, monkeypatch):
    """result_format "packed" / result_uri: the arrays a consumer unpacks are exactly what the default
    JSON form carries (no GPU: the engine is a stand-in, the op's own marshalling is under test)."""
    import json
    import types

    import numpy as np
    from ops import map_classify_tpu as op

    C, k = 6, 3
    rng = np.random.default_rng(0)

    def classify(texts, max_len=128, topk=5, want_logits=False):
        n = len(texts)
        logits = rng.normal(size=(n, C)).astype(np.float32)
        order = np.argsort(-logits, axis=1)[:, :topk].astype(np.int32)
        return {"top_idx": order, "top_score": np.take_along_axis(logits, order, 1), "labels": order[:, 0].copy(),
                "histogram": np.bincount(order[:, 0], minlength=C).astype(np.int64),
                "logits": logits if want_logits else None, "metrics": {"items": n, "gpu_ms_total": 1.0}}

    engine = types.SimpleNamespace(device=0, classify=classify)
    handle = types.SimpleNamespace(interpreter=engine, output_details=[{"shape": (1, C)}], model_path="fake.b2m")
    monkeypatch.setattr(op, "get_cuda_handle", lambda *a, **kw: handle)
    items = [f"item {i}" for i in range(50)]
    rng = np.random.default_rng(0)
    plain = op.run({"items": items, "topk": k, "return_results": True, "return_logits": True, "allow_fallback": False})
    rng = np.random.default_rng(0)
    packed = op.run({"items": items, "topk": k, "result_format": "packed", "return_logits": True,
                     "allow_fallback": False})
    rng = np.random.default_rng(0)
    filed = op.run({"items": items, "topk": k, "result_uri": str(tmp_path / "out"), "allow_fallback": False})
    json.dumps(packed); json.dumps(filed)                                     # still pure JSON (app.py:146)
    assert "labels" not in packed and "labels" not in filed
    assert list(packed.keys())[:4] == ["op", "model_path", "topk", "elapsed_ms"]     # v0 keys stay in front
    assert packed["topk"] == plain["topk"] and packed["label_histogram"] == plain["label_histogram"]
    a = op.unpack_result(packed)
    assert a["labels"].tolist() == plain["labels"] and a["labels"].dtype == np.int32
    assert a["top_index"].tolist() == [[t["index"] for t in r["topk"]] for r in plain["results"]]
    assert np.allclose(a["top_score"], [[t["score"] for t in r["topk"]] for r in plain["results"]])
    assert np.allclose(a["logits"], plain["logits"]) and a["logits"].shape == (50, C)
    f = op.unpack_result(filed)
    assert filed["result_uri"].endswith("out.npz") and filed["result_bytes"] > 0
    assert f["labels"].tolist() == plain["labels"] and f["histogram"].tolist() == plain["label_histogram"]
    assert "logits" not in f
    bad = op.run({"items": items, "result_format": "xml"})
    assert bad["fallback"] == "cpu" and "result_format" in bad["reason"]
