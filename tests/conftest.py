"""Shared fixtures. GPU tests are marked `gpu`; everything else runs on a CPU-only box."""
import hashlib
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "agent-tpu_b200")
ORACLE = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (PKG, ORACLE, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def _cuda_devices() -> int:
    try:
        from ops import _cuda_runtime as rt
        return rt.device_count()
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    """A plain `pytest tests/` on a box without a CUDA device (or without the built library) skips the
    `gpu` tests instead of failing them one by one; `-m gpu` on a GPU box runs them."""
    if not any("gpu" in it.keywords for it in items) or _cuda_devices() > 0:
        return
    skip = pytest.mark.skip(reason="needs a CUDA device and agent-tpu_b200/csrc/libb2classify.so")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def ref_golden():
    with open(os.path.join(GOLDEN, "reference_golden.json"), encoding="utf-8") as f:
        return json.load(f)


@pytest.fixture(scope="session")
def work_dir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("b2work"))


@pytest.fixture(scope="session")
def vocab_path(work_dir):
    import synth
    path = synth.write_vocab(os.path.join(work_dir, "vocab.txt"))
    return path


@pytest.fixture(scope="session")
def vocab_sha(vocab_path):
    return hashlib.sha256(open(vocab_path, "rb").read()).hexdigest()


@pytest.fixture(scope="session")
def tiny_model(work_dir, vocab_path):
    import model_file as mf
    return mf.write_model_file(os.path.join(work_dir, "tiny.b2m"), mf.BERT_TINY)


@pytest.fixture(scope="session")
def base_model(work_dir, vocab_path):
    import model_file as mf
    return mf.write_model_file(os.path.join(work_dir, "base.b2m"), mf.BERT_BASE)


@pytest.fixture(scope="session")
def sentences():
    import synth
    return synth.SentenceGenerator().sentences(2048)


@pytest.fixture()
def clean_env(monkeypatch):
    for k in ("TASKS", "CUDA_MODEL_PATH", "TPU_MODEL_PATH", "CUDA_VOCAB_PATH", "NVIDIA_VISIBLE_DEVICES",
              "TPU_ONLY", "TPU_DISABLED"):
        monkeypatch.delenv(k, raising=False)
    return monkeypatch
