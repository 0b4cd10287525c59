"""The .b2m model file: geometry header plus either f32 tensors or a seed (synthetic weights).

The reference's model is an INT8 `.tflite` file that is not in its repository
(ops/_tpu_runtime.py:27-31), and the contract names no architecture (CONTRACT.md:4-10); SURVEY.md
§8 fixes BERT-base geometry with random-init weights as a builder decision. A synthetic file stores
only the seed: the library regenerates every tensor on the device (csrc/b2_weights.cu) and
`synthetic_tensor` below regenerates the identical f32 values on the host for the oracle, using
integer arithmetic and one multiply-add so the two agree bit for bit.

Layout (little endian) — must match `ModelHeader` in csrc/b2_api.cu:
    char magic[4]="B2MD"; u32 version=1;
    u32 hidden, layers, heads, intermediate, vocab_size, max_positions, type_vocab, num_labels;
    f32 ln_eps; u32 kind (0 tensors follow, 1 synthetic); u64 seed;
    f32 init_std, bias_std, gamma_std, classifier_std;
    [kind 0] f32 tensors in `tensor_specs` order.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, asdict
from typing import Dict, Iterator, List, Tuple

import numpy as np

MAGIC = b"B2MD"
HEADER_FMT = "<4sI8IfIQ4f"
HEADER_SIZE = struct.calcsize(HEADER_FMT)
assert HEADER_SIZE == 72

ROLE_WEIGHT, ROLE_BIAS, ROLE_GAMMA, ROLE_BETA, ROLE_CLASSIFIER = range(5)


@dataclass(frozen=True)
class ModelConfig:
    hidden: int = 768
    layers: int = 12
    heads: int = 12
    intermediate: int = 3072
    vocab_size: int = 30522
    max_positions: int = 512
    type_vocab: int = 2
    num_labels: int = 16
    ln_eps: float = 1e-12
    kind: int = 1
    seed: int = 1234
    init_std: float = 0.02
    bias_std: float = 0.02
    gamma_std: float = 0.05
    classifier_std: float = 0.02


BERT_BASE = ModelConfig()
# small geometry for fast parity tests (same code paths: 2 heads of 64, ragged tile edges)
BERT_TINY = ModelConfig(hidden=128, layers=2, heads=2, intermediate=512, vocab_size=30522,
                        max_positions=512, num_labels=5, init_std=0.08, classifier_std=0.1)


def tensor_specs(cfg: ModelConfig) -> List[Tuple[str, int, Tuple[int, ...], int]]:
    """(name, tensor_id, shape, role) in file order — mirrors tensor_list() in csrc/b2_api.cu."""
    H, I = cfg.hidden, cfg.intermediate
    t = [
        ("word_emb", 0, (cfg.vocab_size, H), ROLE_WEIGHT),
        ("pos_emb", 1, (cfg.max_positions, H), ROLE_WEIGHT),
        ("type_emb", 2, (cfg.type_vocab, H), ROLE_WEIGHT),
        ("emb_ln_g", 3, (H,), ROLE_GAMMA),
        ("emb_ln_b", 4, (H,), ROLE_BETA),
    ]
    for l in range(cfg.layers):
        b = 16 + 16 * l
        t += [
            (f"l{l}.Wqkv", b + 0, (3 * H, H), ROLE_WEIGHT),
            (f"l{l}.bqkv", b + 1, (3 * H,), ROLE_BIAS),
            (f"l{l}.Wo", b + 2, (H, H), ROLE_WEIGHT),
            (f"l{l}.bo", b + 3, (H,), ROLE_BIAS),
            (f"l{l}.ln1_g", b + 4, (H,), ROLE_GAMMA),
            (f"l{l}.ln1_b", b + 5, (H,), ROLE_BETA),
            (f"l{l}.W1", b + 6, (I, H), ROLE_WEIGHT),
            (f"l{l}.b1", b + 7, (I,), ROLE_BIAS),
            (f"l{l}.W2", b + 8, (H, I), ROLE_WEIGHT),
            (f"l{l}.b2", b + 9, (H,), ROLE_BIAS),
            (f"l{l}.ln2_g", b + 10, (H,), ROLE_GAMMA),
            (f"l{l}.ln2_b", b + 11, (H,), ROLE_BETA),
        ]
    t += [
        ("Wp", 8, (H, H), ROLE_WEIGHT),
        ("bp", 9, (H,), ROLE_BIAS),
        ("Wc", 10, (cfg.num_labels, H), ROLE_CLASSIFIER),
        ("bc", 11, (cfg.num_labels,), ROLE_BIAS),
    ]
    return t


_SUM4_STD = np.sqrt(4.0 * (65536.0 * 65536.0 - 1.0) / 12.0)


def _role_scale_shift(cfg: ModelConfig, role: int) -> Tuple[np.float32, np.float32]:
    std, shift = {
        ROLE_WEIGHT: (cfg.init_std, 0.0),
        ROLE_BIAS: (cfg.bias_std, 0.0),
        ROLE_GAMMA: (cfg.gamma_std, 1.0),
        ROLE_BETA: (cfg.bias_std, 0.0),
        ROLE_CLASSIFIER: (cfg.classifier_std, 0.0),
    }[role]
    # the C side computes float(double(float std) / sd): do the same
    return np.float32(np.float64(np.float32(std)) / _SUM4_STD), np.float32(shift)


def synthetic_tensor(seed: int, tensor_id: int, n: int, scale: np.float32, shift: np.float32) -> np.ndarray:
    """Host twin of fill_synth_kernel (csrc/b2_weights.cu): bit-identical f32 values."""
    M = np.uint64(0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        key = np.uint64(seed) ^ (np.uint64(tensor_id) * np.uint64(0xD1B54A32D192ED03) & M)
        out = np.empty(n, dtype=np.float32)
        step = 1 << 22
        for lo in range(0, n, step):
            hi = min(n, lo + step)
            i = np.arange(lo + 1, hi + 1, dtype=np.uint64)
            x = key + i * np.uint64(0x9E3779B97F4A7C15)
            x ^= x >> np.uint64(30)
            x *= np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(27)
            x *= np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(31)
            s = (x & np.uint64(0xFFFF)) + ((x >> np.uint64(16)) & np.uint64(0xFFFF)) + \
                ((x >> np.uint64(32)) & np.uint64(0xFFFF)) + (x >> np.uint64(48))
            f = (s.astype(np.int64) - 131070).astype(np.float32)
            out[lo:hi] = (f * np.float32(scale)).astype(np.float32) + np.float32(shift)
    return out


def write_model_file(path: str, cfg: ModelConfig = BERT_BASE, tensors: Dict[str, np.ndarray] = None) -> str:
    """Write a synthetic (cfg.kind == 1) or explicit-tensor (kind == 0, `tensors` by name) model."""
    kind = 0 if tensors is not None else 1
    hdr = struct.pack(HEADER_FMT, MAGIC, 1, cfg.hidden, cfg.layers, cfg.heads, cfg.intermediate,
                      cfg.vocab_size, cfg.max_positions, cfg.type_vocab, cfg.num_labels, cfg.ln_eps,
                      kind, cfg.seed, cfg.init_std, cfg.bias_std, cfg.gamma_std, cfg.classifier_std)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(hdr)
        if tensors is not None:
            for name, _tid, shape, _role in tensor_specs(cfg):
                a = np.ascontiguousarray(tensors[name], dtype="<f4")
                if tuple(a.shape) != tuple(shape):
                    raise ValueError(f"tensor {name}: shape {a.shape}, expected {shape}")
                f.write(a.tobytes())
    os.replace(tmp, path)
    return path


def read_header(path: str) -> ModelConfig:
    with open(path, "rb") as f:
        raw = f.read(HEADER_SIZE)
    if len(raw) != HEADER_SIZE or raw[:4] != MAGIC:
        raise ValueError(f"{path} is not a B2MD model file")
    (_m, version, hidden, layers, heads, inter, vocab, maxpos, tv, labels, eps, kind, seed,
     istd, bstd, gstd, cstd) = struct.unpack(HEADER_FMT, raw)
    if version != 1:
        raise ValueError(f"{path}: unsupported version {version}")
    return ModelConfig(hidden, layers, heads, inter, vocab, maxpos, tv, labels, eps, kind, seed,
                       istd, bstd, gstd, cstd)


def iter_tensors(path: str) -> Iterator[Tuple[str, np.ndarray]]:
    """Yield (name, f32 array) for every tensor of the model, synthetic or stored."""
    cfg = read_header(path)
    specs = tensor_specs(cfg)
    if cfg.kind == 1:
        for name, tid, shape, role in specs:
            scale, shift = _role_scale_shift(cfg, role)
            n = int(np.prod(shape))
            yield name, synthetic_tensor(cfg.seed, tid, n, scale, shift).reshape(shape)
    else:
        with open(path, "rb") as f:
            f.seek(HEADER_SIZE)
            for name, _tid, shape, _role in specs:
                n = int(np.prod(shape))
                a = np.frombuffer(f.read(4 * n), dtype="<f4")
                if a.size != n:
                    raise ValueError(f"{path}: truncated at tensor {name}")
                yield name, a.reshape(shape).copy()


def load_tensors(path: str) -> Dict[str, np.ndarray]:
    return dict(iter_tensors(path))


def config_dict(cfg: ModelConfig) -> dict:
    return asdict(cfg)
