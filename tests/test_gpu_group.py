"""One job over all GPUs of the box from one process (b2_group_*, ops/_cuda_runtime.EngineGroup).

The reference's agent calls the op once, inline, from one process (app.py:283-287); the group is what
lets that one call use every GPU. Bar: the group's answers are bit-identical to a single device's
(items are independent, slices are contiguous, outputs land in item order), the histogram and the
counters are the all-reduced totals, and a failing job returns an error instead of hanging in the
collective. Runs on however many devices the box has: with one device the slicing and hand-back code
is exercised without NCCL, with two or more (gpurun --gpus N) the NCCL all-reduce as well."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engines(tiny_model, vocab_path):
    from ops import _cuda_runtime as rt
    group = rt.EngineGroup(tiny_model, vocab_path, devices=None, max_batch=64, max_len=128)
    single = rt.Engine(tiny_model, vocab_path, device=0, max_batch=64, max_len=128)
    yield group, single
    group.close()
    single.close()


def test_group_equals_single_device_bit_for_bit(engines, sentences):
    group, single = engines
    from ops import _cuda_runtime as rt
    assert group.n_devices == rt.device_count() and group.devices == list(range(group.n_devices))
    assert (group.nccl_version > 0) == (group.n_devices > 1)
    texts = sentences[:1500] + ["", "x"] + sentences[1500:2041]          # 2,043 items: ragged slices
    before = group.allreduces()
    a = group.classify(texts, 128, 3, want_logits=True)
    b = single.classify(texts, 128, 3, want_logits=True)
    for key in ("logits", "top_idx", "top_score", "labels", "histogram"):
        assert (a[key] == b[key]).all(), key
    m = a["metrics"]
    assert m["items"] == len(texts) and m["tokens"] == b["metrics"]["tokens"]
    assert m["devices"] == group.devices and len(m["per_device"]) == group.n_devices
    W = group.n_devices
    assert [d["items"] for d in m["per_device"]] == [(i + 1) * len(texts) // W - i * len(texts) // W for i in range(W)]
    assert sum(d["tokens"] for d in m["per_device"]) == m["tokens"]
    assert abs(m["gpu_ms_total"] - max(d["gpu_ms_total"] for d in m["per_device"])) < 1e-6 * max(1.0, m["gpu_ms_total"])
    assert group.allreduces() - before == (2 if W > 1 else 0)


def test_group_jobs_smaller_than_the_group(engines, sentences):
    group, single = engines
    for n in (0, 1, 2, group.n_devices, group.n_devices + 1):
        a = group.classify(sentences[:n], 128, 2, want_logits=True)
        b = single.classify(sentences[:n], 128, 2, want_logits=True)
        assert a["logits"].shape == (n, 5)
        assert (a["logits"] == b["logits"]).all() and (a["histogram"] == b["histogram"]).all()
        assert a["metrics"]["items"] == n


def test_group_failure_is_an_error_on_the_caller_not_a_hang(engines, sentences):
    from ops import _cuda_runtime as rt
    group, single = engines
    with pytest.raises(rt.B2Error) as e:
        group.classify(sentences[:100], 100, 2)              # max_len not a multiple of 128: every slice refuses
    assert "device" in str(e.value) and "multiple of 128" in str(e.value)
    with pytest.raises(rt.B2Error):
        group.classify(sentences[:100], 128, 99)             # topk beyond the number of labels
    out = group.classify(sentences[:100], 128, 2, want_logits=True)   # and the group still works
    assert (out["logits"] == single.classify(sentences[:100], 128, 2, want_logits=True)["logits"]).all()


def test_group_csv_chain_equals_single_device(engines, sentences):
    """BASELINE config 5's job shape: file -> index (device 0) -> text column slices sent device to
    device -> tokenizer -> encoder on every device -> NCCL histogram reduce."""
    from ops import _cuda_runtime as rt
    group, single = engines
    rows = sentences[:700]
    table = ("id,text,source\r\n" + "".join(
        f'{i},"{t.replace(chr(34), chr(34) * 2)}",s{i % 3}\r\n' for i, t in enumerate(rows))).encode()
    gi = rt.CsvIndex(group, table)
    si = rt.CsvIndex(single, table)
    try:
        assert gi.rows == si.rows == len(rows)
        for start, count in ((0, len(rows)), (13, 501), (690, 10), (5, 1), (7, 0)):
            a = gi.classify(1, start, count, 128, 3, want_logits=True)
            b = si.classify(1, start, count, 128, 3, want_logits=True)
            for key in ("logits", "top_idx", "labels", "histogram"):
                assert (a[key] == b[key]).all(), (key, start, count)
            assert a["metrics"]["items"] == count and a["metrics"]["tokens"] == b["metrics"]["tokens"]
        plain = single.classify(rows, 128, 3, want_logits=True)
        assert (gi.classify(1, 0, len(rows), 128, 3, want_logits=True)["logits"] == plain["logits"]).all()
    finally:
        gi.close(); si.close()


def test_op_uses_every_device_of_the_box(tiny_model, vocab_path, clean_env, sentences, tmp_path):
    """ops.get_op("map_classify_tpu")(payload), called once from one process, spreads the job over
    all devices (B2_DEVICES narrows it); under a one-process-per-GPU launcher it keeps to its own."""
    import ops
    from ops import _cuda_runtime as rt
    rt.reset_cache()
    clean_env.delenv("B2_DEVICES", raising=False)
    clean_env.delenv("B2_DEVICE", raising=False)
    clean_env.delenv("LOCAL_RANK", raising=False)
    clean_env.setenv("CUDA_MODEL_PATH", tiny_model)
    clean_env.setenv("CUDA_VOCAB_PATH", vocab_path)
    n_dev = rt.device_count()
    assert rt.job_devices() == list(range(n_dev))
    run = ops.get_op("map_classify_tpu")
    texts = sentences[:333]
    r = run({"items": texts, "topk": 2, "batch": 32, "allow_fallback": False})
    eng = rt.get_cuda_handle(tiny_model, vocab_path, max_len=128, max_batch=32).interpreter
    assert isinstance(eng, rt.EngineGroup) == (n_dev > 1)
    if n_dev > 1:
        assert r["metrics"]["devices"] == list(range(n_dev)) and len(r["metrics"]["per_device"]) == n_dev
    one = rt.Engine(tiny_model, vocab_path, device=0, max_batch=32, max_len=128)
    try:
        want = one.classify(texts, 128, 2)
    finally:
        one.close()
    assert r["labels"] == want["labels"].tolist() and r["label_histogram"] == want["histogram"].tolist()
    assert r["items_count"] == 333 and r["metrics"]["items"] == 333
    # the chained CSV job through the op
    path = tmp_path / "t.csv"
    path.write_text("id,text\n" + "".join(f'{i},"{t}"\n' for i, t in enumerate(texts)), encoding="utf-8")
    c = run({"source_uri": str(path), "text_column": "text", "topk": 2, "batch": 32, "allow_fallback": False})
    assert c["labels"] == r["labels"] and c["label_histogram"] == r["label_histogram"] and c["row_count"] == 333
    rt.reset_cache()
    clean_env.setenv("B2_DEVICES", "0")
    assert rt.job_devices() == [0]
    clean_env.setenv("LOCAL_RANK", str(n_dev - 1))
    assert rt.job_devices() == [n_dev - 1]
    rt.reset_cache()
