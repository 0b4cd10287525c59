"""B200 accelerator runtime for the ops: replaces the reference's ops/_tpu_runtime.py.

Same shape as the reference module (ops/_tpu_runtime.py:15-63):
  * `get_model_path(requested)`  — precedence: payload, environment, built-in default;
  * `get_cuda_handle(model_path)` — process-wide single-slot cache keyed on the path, lazy import of
    the native dependency so the agent starts without it, device probe, file-existence check, load;
  * a frozen handle dataclass `(interpreter, input_details, output_details, model_path)`.

Underneath, the "interpreter" is `Engine`, a ctypes binding of libb2classify.so
(include/b2_classify.h). There is no CPU fallback: if the library or a CUDA device is missing the
calls raise, and the op's own `except` turns that into the reference's error envelope
(ops/map_classify_tpu.py:84-90).
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# $B2_LIB_PATH: another build of the same library (A/B runs of two kernel versions on one GPU box)
LIB_PATH = os.environ.get("B2_LIB_PATH") or os.path.join(_PKG_DIR, "csrc", "libb2classify.so")
UNICODE_TABLE_PATH = os.path.join(_PKG_DIR, "data", "bert_unicode.bin")
DEFAULT_MODEL_PATH = "/models/model_b200.b2m"

# Lazy, process-wide cache (single slot, like ops/_tpu_runtime.py:9-12)
_ENGINE = None
_INPUT_DETAILS = None
_OUTPUT_DETAILS = None
_MODEL_PATH = None
_CACHE_KEY = None
_LOCK = threading.Lock()  # the reference's cache is unlocked; creation is locked here
_LIB = None


class B2Error(RuntimeError):
    """A non-zero return code from libb2classify.so."""

    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


class _Config(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("device", C.c_int32), ("model_path", C.c_char_p),
                ("vocab_path", C.c_char_p), ("unicode_path", C.c_char_p), ("max_batch", C.c_int32),
                ("max_len", C.c_int32)]


class ModelInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("hidden", "layers", "heads", "intermediate", "vocab_size",
                                         "max_positions", "num_labels", "tokenizer_vocab_size",
                                         "max_batch", "max_len", "device", "sm_count")] + \
               [("weight_bytes", C.c_int64), ("activation_bytes", C.c_int64), ("packed_rows", C.c_int32),
                ("ln_folded", C.c_int32)]


_METRIC_KINDS = ("tokenize", "embed", "gemm_qkv", "attention", "gemm_attn_out", "layernorm", "gemm_ffn1",
                 "gemm_ffn2", "pooler", "head")


class Metrics(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("profiled", C.c_int32)] + \
               [(n, C.c_int64) for n in ("items", "tokens", "padded_tokens", "batches", "kernel_launches",
                                         "h2d_bytes", "d2h_bytes")] + \
               [("gpu_ms_total", C.c_double)] + [("ms_" + k, C.c_double) for k in _METRIC_KINDS] + \
               [("n_" + k, C.c_int64) for k in _METRIC_KINDS]

    def as_dict(self) -> dict:
        d = {n: int(getattr(self, n)) for n in ("items", "tokens", "padded_tokens", "batches",
                                                 "kernel_launches", "h2d_bytes", "d2h_bytes")}
        d["gpu_ms_total"] = float(self.gpu_ms_total)
        if self.profiled:
            d["kernels"] = {k: {"ms": float(getattr(self, "ms_" + k)), "launches": int(getattr(self, "n_" + k))}
                            for k in _METRIC_KINDS}
        return d


class DeviceInfo(C.Structure):
    _fields_ = [("name", C.c_char * 128), ("total_memory_bytes", C.c_int64), ("sm_count", C.c_int32),
                ("cc_major", C.c_int32), ("cc_minor", C.c_int32), ("l2_bytes", C.c_int32)]


class CsvInfo(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("has_header", C.c_int32), ("header_fields", C.c_int32),
                ("end_state", C.c_int32), ("header_bytes", C.c_int64), ("n_bytes", C.c_int64),
                ("data_offset", C.c_int64), ("rows", C.c_int64), ("first_invalid_utf8", C.c_int64),
                ("header_field_too_long", C.c_int32), ("reserved", C.c_int32),
                ("header_long_offset", C.c_int64), ("long_field_row", C.c_int64),
                ("long_field_offset", C.c_int64), ("gpu_ms", C.c_double), ("kernel_launches", C.c_int64),
                ("index_ms", C.c_double)]


def load_library():
    """dlopen libb2classify.so and declare its prototypes. Raises OSError if it is not built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise OSError(f"libb2classify.so is not built: {LIB_PATH} (run `python -c 'import __graft_entry__ "
                      f"as g; g.build()'` or `make -C agent-tpu_b200/csrc`)")
    lib = C.CDLL(LIB_PATH)
    P, I32, I64 = C.c_void_p, C.c_int32, C.c_int64
    lib.b2_last_error.restype = C.c_char_p
    lib.b2_last_error.argtypes = []
    lib.b2_device_count.restype = I32
    lib.b2_device_count.argtypes = []
    lib.b2_get_device_info.argtypes = [I32, C.POINTER(DeviceInfo)]
    lib.b2_get_device_memory.argtypes = [I32, C.POINTER(I64), C.POINTER(I64)]
    lib.b2_get_device_memory.restype = C.c_int
    lib.b2_get_peer_info.argtypes = [I32, I32, C.POINTER(I32), C.POINTER(I32), C.POINTER(I32)]
    lib.b2_get_peer_info.restype = C.c_int
    lib.b2_create.argtypes = [C.POINTER(_Config), C.POINTER(P)]
    lib.b2_destroy.argtypes = [P]
    lib.b2_destroy.restype = None
    lib.b2_get_model_info.argtypes = [P, C.POINTER(ModelInfo)]
    lib.b2_set_profiling.argtypes = [P, I32]
    lib.b2_get_stream.argtypes = [P]
    lib.b2_get_stream.restype = C.c_uint64
    lib.b2_tokenize.argtypes = [P, P, P, I64, I32, P, P, C.POINTER(Metrics)]
    lib.b2_tokenize_device.argtypes = [P, P, P, I64, I64, I32, P, P, C.POINTER(Metrics)]
    lib.b2_classify_text.argtypes = [P, P, P, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics)]
    lib.b2_classify_text_device.argtypes = [P, P, P, I64, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics)]
    lib.b2_classify_ids.argtypes = [P, P, P, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics)]
    lib.b2_chunk_count.argtypes = [P, P, P, I64, P]
    lib.b2_chunk_bounds.argtypes = [P, P, P, I64, I64, P, I64, P]
    lib.b2_csv_open.argtypes = [P, P, I64, C.POINTER(P), C.POINTER(CsvInfo)]
    lib.b2_csv_close.argtypes = [P]
    lib.b2_csv_close.restype = None
    lib.b2_csv_header.argtypes = [P, P, I64, P, I32]
    lib.b2_csv_row_spans.argtypes = [P, I64, I64, P, P]
    lib.b2_csv_rows_size.argtypes = [P, I64, I64, I32, C.POINTER(I64), C.POINTER(I64)]
    lib.b2_csv_rows_read.argtypes = [P, I64, I64, I32, P, P, P]
    lib.b2_csv_classify.argtypes = [P, I32, I64, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics)]
    lib.b2_group_create.argtypes = [C.POINTER(_Config), P, I32, C.POINTER(P)]
    lib.b2_group_destroy.argtypes = [P]
    lib.b2_group_destroy.restype = None
    lib.b2_group_size.argtypes = [P]
    lib.b2_group_size.restype = I32
    lib.b2_group_member.argtypes = [P, I32]
    lib.b2_group_member.restype = P
    lib.b2_group_info.argtypes = [P, P, I32, C.POINTER(I32), C.POINTER(I64)]
    lib.b2_group_classify_text.argtypes = [P, P, P, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics), P]
    lib.b2_group_csv_classify.argtypes = [P, P, I32, I64, I64, I32, I32, P, P, P, P, P, C.POINTER(Metrics), P]
    for name in ("b2_group_create", "b2_group_info", "b2_group_classify_text", "b2_group_csv_classify"):
        getattr(lib, name).restype = C.c_int
    lib.b2_tokenize_noalloc.argtypes = [P, P, P, I64, I32, P, P]
    lib.b2_tokenize_noalloc.restype = C.c_int
    lib.b2_debug_hidden_states.argtypes = [P, P, P, I64, I32, P, P]
    lib.b2_debug_hidden_states.restype = C.c_int
    for name in ("b2_csv_open", "b2_csv_header", "b2_csv_row_spans", "b2_csv_rows_size", "b2_csv_rows_read",
                 "b2_csv_classify"):
        getattr(lib, name).restype = C.c_int
    for name in ("b2_get_device_info", "b2_create", "b2_get_model_info", "b2_set_profiling", "b2_tokenize",
                 "b2_tokenize_device", "b2_classify_text", "b2_classify_text_device", "b2_classify_ids",
                 "b2_chunk_count", "b2_chunk_bounds"):
        getattr(lib, name).restype = C.c_int
    _LIB = lib
    return lib


def _check(lib, rc: int) -> None:
    if rc != 0:
        msg = lib.b2_last_error()
        raise B2Error(rc, (msg or b"").decode("utf-8", "replace") or f"libb2classify error {rc}")


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_items(items: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """Concatenate items as UTF-8 (uint8 array) and int64 offsets[n+1] — the C ABI's text layout.

    The boundary of Python at 1M items is one join and one cumsum, not a per-item ctypes call
    (SURVEY.md §7.2 "Python at the boundary dominates")."""
    n = len(items)
    offsets = np.zeros(n + 1, dtype=np.int64)
    if not n:
        return np.zeros(0, dtype=np.uint8), offsets
    joined = "".join(items)
    if joined.isascii():  # one encode for the whole job; byte lengths are the character lengths
        data = joined.encode("ascii")
        np.cumsum(np.fromiter(map(len, items), dtype=np.int64, count=n), out=offsets[1:])
    else:
        enc = [s.encode("utf-8", "surrogatepass") for s in items]
        np.cumsum(np.fromiter(map(len, enc), dtype=np.int64, count=n), out=offsets[1:])
        data = b"".join(enc)
    blob = np.frombuffer(data, dtype=np.uint8) if data else np.zeros(0, dtype=np.uint8)
    return blob, offsets


class Engine:
    """One model + tokenizer resident on one CUDA device (the reference's `interpreter` slot)."""

    def __init__(self, model_path: Optional[str], vocab_path: Optional[str], device: int = 0,
                 max_batch: int = 1024, max_len: int = 128):
        self._lib = load_library()
        self._h = C.c_void_p()
        cfg = _Config(C.sizeof(_Config), int(device),
                      model_path.encode() if model_path else None,
                      vocab_path.encode() if vocab_path else None,
                      UNICODE_TABLE_PATH.encode() if vocab_path else None,
                      int(max_batch), int(max_len))
        _check(self._lib, self._lib.b2_create(C.byref(cfg), C.byref(self._h)))
        self.info = ModelInfo()
        _check(self._lib, self._lib.b2_get_model_info(self._h, C.byref(self.info)))
        self.model_path = model_path
        self.vocab_path = vocab_path
        self.device = int(device)

    # -- life cycle --------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.b2_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    @property
    def stream(self) -> int:
        """cudaStream_t (as int) that this engine's kernels run on."""
        return int(self._lib.b2_get_stream(self._h))

    def set_profiling(self, on: bool) -> None:
        _check(self._lib, self._lib.b2_set_profiling(self._h, 1 if on else 0))

    # -- tokenizer ---------------------------------------------------------------------------
    def tokenize_packed(self, blob: np.ndarray, offsets: np.ndarray, max_len: int,
                        want_mask: bool = True) -> Tuple[np.ndarray, Optional[np.ndarray], dict]:
        n = len(offsets) - 1
        ids = np.empty((n, max_len), dtype=np.int32)
        mask = np.empty((n, max_len), dtype=np.uint8) if want_mask else None
        m = Metrics(struct_size=C.sizeof(Metrics))
        _check(self._lib, self._lib.b2_tokenize(self._h, _ptr(blob), _ptr(offsets), n, int(max_len),
                                                _ptr(ids), _ptr(mask), C.byref(m)))
        return ids, mask, m.as_dict()

    def tokenize(self, items: Sequence[str], max_len: int = 128):
        blob, offsets = pack_items(items)
        return self.tokenize_packed(blob, offsets, max_len)

    # -- classification ----------------------------------------------------------------------
    def _outputs(self, n: int, topk: int, want_logits: bool):
        Cn = int(self.info.num_labels)
        return (np.empty((n, topk), dtype=np.int32), np.empty((n, topk), dtype=np.float32),
                np.empty(n, dtype=np.int32), np.zeros(Cn, dtype=np.int64),
                np.empty((n, Cn), dtype=np.float32) if want_logits else None)

    def classify_packed(self, blob: np.ndarray, offsets: np.ndarray, max_len: int = 128, topk: int = 5,
                        want_logits: bool = False, outputs=None) -> dict:
        """`outputs` may supply (top_idx, top_score, labels, hist, logits) arrays to fill, e.g. views
        of pinned host memory; otherwise fresh numpy arrays are returned."""
        n = len(offsets) - 1
        top_idx, top_score, labels, hist, logits = outputs or self._outputs(n, topk, want_logits)
        m = Metrics(struct_size=C.sizeof(Metrics))
        _check(self._lib, self._lib.b2_classify_text(self._h, _ptr(blob), _ptr(offsets), n, int(max_len),
                                                     int(topk), _ptr(top_idx), _ptr(top_score), _ptr(labels),
                                                     _ptr(hist), _ptr(logits), C.byref(m)))
        return {"top_idx": top_idx, "top_score": top_score, "labels": labels, "histogram": hist,
                "logits": logits, "metrics": m.as_dict()}

    def classify(self, items: Sequence[str], max_len: int = 128, topk: int = 5,
                 want_logits: bool = False) -> dict:
        blob, offsets = pack_items(items)
        return self.classify_packed(blob, offsets, max_len, topk, want_logits)

    def classify_ids(self, ids: np.ndarray, mask: np.ndarray, topk: int = 5, want_logits: bool = False) -> dict:
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        if ids.ndim != 2 or mask.shape != ids.shape:
            raise ValueError("ids and mask must be [n, max_len] arrays of the same shape")
        n, max_len = ids.shape
        top_idx, top_score, labels, hist, logits = self._outputs(n, topk, want_logits)
        m = Metrics(struct_size=C.sizeof(Metrics))
        _check(self._lib, self._lib.b2_classify_ids(self._h, _ptr(ids), _ptr(mask), n, int(max_len), int(topk),
                                                    _ptr(top_idx), _ptr(top_score), _ptr(labels), _ptr(hist),
                                                    _ptr(logits), C.byref(m)))
        return {"top_idx": top_idx, "top_score": top_score, "labels": labels, "histogram": hist,
                "logits": logits, "metrics": m.as_dict()}

    def debug_hidden_states(self, ids: np.ndarray, mask: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """Test aid (b2_debug_hidden_states): one batch of pre-tokenised rows through the production
        kernels; returns (hidden [layers + 1, n, max_len, H] float32, logits [n, num_labels])."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        if ids.ndim != 2 or mask.shape != ids.shape:
            raise ValueError("ids and mask must be [n, max_len] arrays of the same shape")
        n, max_len = ids.shape
        hidden = np.empty((int(self.info.layers) + 1, n, max_len, int(self.info.hidden)), dtype=np.float32)
        logits = np.empty((n, int(self.info.num_labels)), dtype=np.float32)
        _check(self._lib, self._lib.b2_debug_hidden_states(self._h, _ptr(ids), _ptr(mask), n, int(max_len),
                                                           _ptr(hidden), _ptr(logits)))
        return hidden, logits

    def tokenize_noalloc(self, blob: np.ndarray, offsets: np.ndarray, max_len: int,
                         want_mask: bool = True) -> Tuple[np.ndarray, Optional[np.ndarray]]:
        """b2_tokenize_noalloc: same answers as tokenize_packed, staged in the library's fixed 8 MiB
        device block instead of buffers that may have to grow — for callers that must not allocate
        device memory while running. Jobs beyond the block raise B2Error (code 5)."""
        n = len(offsets) - 1
        ids = np.empty((n, max_len), dtype=np.int32)
        mask = np.empty((n, max_len), dtype=np.uint8) if want_mask else None
        _check(self._lib, self._lib.b2_tokenize_noalloc(self._h, _ptr(blob), _ptr(offsets), n, int(max_len),
                                                        _ptr(ids), _ptr(mask)))
        return ids, mask

    def classify_device(self, d_text: int, d_offsets: int, n: int, n_bytes: int, max_len: int, topk: int,
                        d_top_idx: int, d_top_score: int, d_labels: int, d_hist: int, d_logits: int = 0) -> dict:
        """Everything already resident in this device's HBM (raw device addresses)."""
        m = Metrics(struct_size=C.sizeof(Metrics))
        _check(self._lib, self._lib.b2_classify_text_device(
            self._h, C.c_void_p(d_text), C.c_void_p(d_offsets), int(n), int(n_bytes), int(max_len), int(topk),
            C.c_void_p(d_top_idx), C.c_void_p(d_top_score), C.c_void_p(d_labels or None),
            C.c_void_p(d_hist or None), C.c_void_p(d_logits or None), C.byref(m)))
        return m.as_dict()

    def tokenize_device(self, d_text: int, d_offsets: int, n: int, n_bytes: int, max_len: int,
                        d_ids: int, d_mask: int = 0) -> dict:
        m = Metrics(struct_size=C.sizeof(Metrics))
        _check(self._lib, self._lib.b2_tokenize_device(
            self._h, C.c_void_p(d_text), C.c_void_p(d_offsets), int(n), int(n_bytes), int(max_len),
            C.c_void_p(d_ids), C.c_void_p(d_mask or None), C.byref(m)))
        return m.as_dict()

    # -- fixed-width chunking (map_tokenize) -------------------------------------------------
    def chunk_counts(self, blob: np.ndarray, offsets: np.ndarray) -> np.ndarray:
        n = len(offsets) - 1
        counts = np.zeros(n, dtype=np.int64)
        _check(self._lib, self._lib.b2_chunk_count(self._h, _ptr(blob), _ptr(offsets), n, _ptr(counts)))
        return counts

    def chunk_bounds(self, blob: np.ndarray, offsets: np.ndarray, chunk_size: int,
                     chunk_base: np.ndarray, n_chunks: int) -> np.ndarray:
        n = len(offsets) - 1
        bounds = np.zeros(max(n_chunks, 1), dtype=np.int64)
        base = np.ascontiguousarray(chunk_base, dtype=np.int64)
        _check(self._lib, self._lib.b2_chunk_bounds(self._h, _ptr(blob), _ptr(offsets), n, int(chunk_size),
                                                    _ptr(base), int(n_chunks), _ptr(bounds)))
        return bounds[:n_chunks]


def _hint_nccl() -> None:
    """Point the library at the libnccl.so.2 that PyTorch ships when the process has not loaded one yet
    and the caller has not chosen one ($B2_NCCL_PATH): if torch is imported later it then finds its own
    copy already in place instead of a second, different one."""
    if os.environ.get("B2_NCCL_PATH"):
        return
    import sys
    for base in sys.path:
        cand = os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["B2_NCCL_PATH"] = cand
            return


class EngineGroup(Engine):
    """One model + tokenizer resident on several GPUs of this box, driven from this one process
    (b2_group_*): a job's items are cut into contiguous slices, one per device, and the label
    histogram and counters are summed with one NCCL all-reduce per job. Everything an Engine can do
    still works and runs on the first device (`handle` is member 0's); `classify*` use all of them."""

    def __init__(self, model_path: str, vocab_path: Optional[str], devices: Optional[Sequence[int]] = None,
                 max_batch: int = 1024, max_len: int = 128):
        self._lib = load_library()
        self._g = C.c_void_p()
        self._h = C.c_void_p()
        devs = [int(d) for d in devices] if devices else []
        if len(devs) != 1:
            _hint_nccl()
        cfg = _Config(C.sizeof(_Config), 0, model_path.encode() if model_path else None,
                      vocab_path.encode() if vocab_path else None,
                      UNICODE_TABLE_PATH.encode() if vocab_path else None, int(max_batch), int(max_len))
        arr = (C.c_int32 * len(devs))(*devs) if devs else None
        _check(self._lib, self._lib.b2_group_create(C.byref(cfg), arr, len(devs), C.byref(self._g)))
        self.n_devices = int(self._lib.b2_group_size(self._g))
        self._h = C.c_void_p(self._lib.b2_group_member(self._g, 0))
        self.info = ModelInfo()
        _check(self._lib, self._lib.b2_get_model_info(self._h, C.byref(self.info)))
        out = (C.c_int32 * self.n_devices)()
        ver = C.c_int32(0)
        _check(self._lib, self._lib.b2_group_info(self._g, out, self.n_devices, C.byref(ver), None))
        self.devices = [int(d) for d in out]
        self.nccl_version = int(ver.value)
        self.model_path = model_path
        self.vocab_path = vocab_path
        self.device = self.devices[0]

    def close(self) -> None:
        if getattr(self, "_g", None) is not None and self._g.value:
            self._lib.b2_group_destroy(self._g)     # destroys the member handles too
            self._g = C.c_void_p()
            self._h = C.c_void_p()

    @property
    def group(self) -> C.c_void_p:
        return self._g

    def member(self, i: int) -> C.c_void_p:
        return C.c_void_p(self._lib.b2_group_member(self._g, int(i)))

    def allreduces(self) -> int:
        n = C.c_int64(0)
        _check(self._lib, self._lib.b2_group_info(self._g, None, 0, None, C.byref(n)))
        return int(n.value)

    def set_profiling(self, on: bool) -> None:
        for i in range(self.n_devices):
            _check(self._lib, self._lib.b2_set_profiling(self.member(i), 1 if on else 0))

    def _group_metrics(self, total: "Metrics", per) -> dict:
        d = total.as_dict()
        d["devices"] = list(self.devices)
        d["per_device"] = [m.as_dict() for m in per]
        return d

    def classify_packed(self, blob: np.ndarray, offsets: np.ndarray, max_len: int = 128, topk: int = 5,
                        want_logits: bool = False, outputs=None) -> dict:
        n = len(offsets) - 1
        top_idx, top_score, labels, hist, logits = outputs or self._outputs(n, topk, want_logits)
        m = Metrics(struct_size=C.sizeof(Metrics))
        per = (Metrics * self.n_devices)()
        for p in per:
            p.struct_size = C.sizeof(Metrics)
        _check(self._lib, self._lib.b2_group_classify_text(self._g, _ptr(blob), _ptr(offsets), n, int(max_len),
                                                           int(topk), _ptr(top_idx), _ptr(top_score), _ptr(labels),
                                                           _ptr(hist), _ptr(logits), C.byref(m), per))
        return {"top_idx": top_idx, "top_score": top_score, "labels": labels, "histogram": hist,
                "logits": logits, "metrics": self._group_metrics(m, per)}


class CsvIndex:
    """A CSV file resident in one engine's HBM with its record index (b2_csv_open): the device-side
    stand-in for the `csv.DictReader` loop of _read_csv_shard (ops/csv_shard.py:9-26)."""

    def __init__(self, engine: Engine, data):
        self._lib = engine._lib
        self._engine = engine  # keeps the handle alive
        self._c = C.c_void_p()
        buf = np.frombuffer(data, dtype=np.uint8) if len(data) else np.zeros(0, dtype=np.uint8)
        self.info = CsvInfo(struct_size=C.sizeof(CsvInfo))
        _check(self._lib, self._lib.b2_csv_open(engine.handle, _ptr(buf) if len(buf) else None, len(buf),
                                                C.byref(self._c), C.byref(self.info)))
        self.rows = int(self.info.rows)
        self.n_bytes = int(self.info.n_bytes)

    def close(self) -> None:
        if getattr(self, "_c", None) is not None and self._c.value:
            if self._engine.handle.value:  # closing the engine first frees the device context's memory
                self._lib.b2_csv_close(self._c)
            self._c = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def fieldnames(self) -> Optional[List[str]]:
        """DictReader.fieldnames: None when the data holds no record at all."""
        if not self.info.has_header:
            return None
        k = int(self.info.header_fields)
        names = np.zeros(max(int(self.info.header_bytes), 1), dtype=np.uint8)
        ends = np.zeros(max(k, 1), dtype=np.int64)
        _check(self._lib, self._lib.b2_csv_header(self._c, _ptr(names), len(names), _ptr(ends), len(ends)))
        raw = names.tobytes()
        out, a = [], 0
        for e in ends[:k].tolist():
            out.append(raw[a:e].decode("utf-8"))
            a = e
        return out

    def row_spans(self, start_row: int, n_rows: int) -> Tuple[np.ndarray, np.ndarray]:
        a = np.zeros(max(n_rows, 1), dtype=np.int64)
        e = np.zeros(max(n_rows, 1), dtype=np.int64)
        _check(self._lib, self._lib.b2_csv_row_spans(self._c, int(start_row), int(n_rows), _ptr(a), _ptr(e)))
        return a[:n_rows], e[:n_rows]

    def rows_read(self, start_row: int, n_rows: int, column: int = -1):
        """(content bytes, field_end[n_fields], row_field_end[n_rows]) of the unquoted fields."""
        nf, nb = C.c_int64(0), C.c_int64(0)
        _check(self._lib, self._lib.b2_csv_rows_size(self._c, int(start_row), int(n_rows), int(column),
                                                     C.byref(nf), C.byref(nb)))
        blob = np.zeros(max(nb.value, 1), dtype=np.uint8)
        fend = np.zeros(max(nf.value, 1), dtype=np.int64)
        rend = np.zeros(max(n_rows, 1), dtype=np.int64)
        _check(self._lib, self._lib.b2_csv_rows_read(self._c, int(start_row), int(n_rows), int(column),
                                                     _ptr(blob), _ptr(fend), _ptr(rend)))
        return blob[:nb.value], fend[:nf.value], rend[:n_rows]

    def classify(self, column: int, start_row: int, n_rows: int, max_len: int = 128, topk: int = 5,
                 want_logits: bool = False) -> dict:
        """csv -> tokenize -> classify without the text leaving the device (b2_csv_classify)."""
        top_idx, top_score, labels, hist, logits = self._engine._outputs(n_rows, topk, want_logits)
        m = Metrics(struct_size=C.sizeof(Metrics))
        eng = self._engine
        if isinstance(eng, EngineGroup):
            # the file lives on the group's first device; slices of the text column travel to the
            # other devices peer to peer (b2_group_csv_classify)
            per = (Metrics * eng.n_devices)()
            for p in per:
                p.struct_size = C.sizeof(Metrics)
            _check(self._lib, self._lib.b2_group_csv_classify(eng.group, self._c, int(column), int(start_row),
                                                              int(n_rows), int(max_len), int(topk), _ptr(top_idx),
                                                              _ptr(top_score), _ptr(labels), _ptr(hist),
                                                              _ptr(logits), C.byref(m), per))
            metrics = eng._group_metrics(m, per)
        else:
            _check(self._lib, self._lib.b2_csv_classify(self._c, int(column), int(start_row), int(n_rows),
                                                        int(max_len), int(topk), _ptr(top_idx), _ptr(top_score),
                                                        _ptr(labels), _ptr(hist), _ptr(logits), C.byref(m)))
            metrics = m.as_dict()
        return {"top_idx": top_idx, "top_score": top_score, "labels": labels, "histogram": hist,
                "logits": logits, "metrics": metrics}


@dataclass(frozen=True)
class CUDAHandle:
    """Counterpart of TPUHandle (ops/_tpu_runtime.py:15-20)."""
    interpreter: object      # Engine
    input_details: object    # [{"index", "shape", "dtype"}] — one pre-tokenised row, like the v0 contract
    output_details: object   # [{"index", "shape", "dtype"}] — one row of class scores
    model_path: str


def get_model_path(requested: Optional[str] = None) -> str:
    """Payload value, else $CUDA_MODEL_PATH, else $TPU_MODEL_PATH (kept so an existing deployment's
    environment still works), else the default — same precedence as ops/_tpu_runtime.py:23-31."""
    return (
        requested
        or os.environ.get("CUDA_MODEL_PATH")
        or os.environ.get("TPU_MODEL_PATH")
        or DEFAULT_MODEL_PATH
    )


def get_vocab_path(model_path: str, requested: Optional[str] = None) -> str:
    """vocab.txt: payload value, else $CUDA_VOCAB_PATH, else `vocab.txt` beside the model file."""
    return (
        requested
        or os.environ.get("CUDA_VOCAB_PATH")
        or os.path.join(os.path.dirname(os.path.abspath(model_path)), "vocab.txt")
    )


def device_count() -> int:
    return int(load_library().b2_device_count())


def device_infos() -> List[dict]:
    """Per-device facts for worker_sizing (replaces the nvidia-smi subprocess)."""
    lib = load_library()
    out = []
    for i in range(int(lib.b2_device_count())):
        d = DeviceInfo()
        _check(lib, lib.b2_get_device_info(i, C.byref(d)))
        entry = {"index": i, "name": d.name.decode("utf-8", "replace"),
                 "total_memory_bytes": int(d.total_memory_bytes), "sm_count": int(d.sm_count),
                 "compute_capability": f"{d.cc_major}.{d.cc_minor}", "l2_bytes": int(d.l2_bytes)}
        try:  # what a controller needs to size `batch` x `max_len` for this worker right now
            free, _total = device_memory(i)
            entry["free_memory_bytes"] = free
            entry["classify_limits"] = classify_limits(free)
        except B2Error:
            pass
        out.append(entry)
    return out


def interconnect() -> dict:
    """How the box's GPUs reach each other's memory: the peer matrix (direct access, the driver's link
    rank, atomics) and two summaries — every pair directly addressable, and every pair with native
    atomics, which on this hardware means NVLink / NVSwitch rather than PCIe."""
    m = peer_matrix()
    n = len(m)
    pairs = [(a, b) for a in range(n) for b in range(n) if a != b]
    return {"gpu_count": n,
            "peer_access": [[bool(m[a][b]["access"]) for b in range(n)] for a in range(n)],
            "performance_rank": [[int(m[a][b]["rank"]) for b in range(n)] for a in range(n)],
            "all_to_all_peer_access": bool(pairs) and all(m[a][b]["access"] for a, b in pairs),
            "nvlink_all_to_all": bool(pairs) and all(m[a][b]["access"] and m[a][b]["atomics"] for a, b in pairs)}


def device_memory(device: int = 0) -> Tuple[int, int]:
    """(free, total) bytes of HBM on `device` right now."""
    lib = load_library()
    f, t = C.c_int64(0), C.c_int64(0)
    _check(lib, lib.b2_get_device_memory(int(device), C.byref(f), C.byref(t)))
    return int(f.value), int(t.value)


def peer_matrix() -> List[List[dict]]:
    """peer_matrix()[a][b] = {"access", "rank", "atomics"}: how device a reaches device b's memory."""
    lib = load_library()
    n = int(lib.b2_device_count())
    out = []
    for a in range(n):
        row = []
        for b in range(n):
            ca, pr, na = C.c_int32(0), C.c_int32(0), C.c_int32(0)
            _check(lib, lib.b2_get_peer_info(a, b, C.byref(ca), C.byref(pr), C.byref(na)))
            row.append({"access": bool(ca.value), "rank": int(pr.value), "atomics": bool(na.value)})
        out.append(row)
    return out


def activation_bytes_per_row(hidden: int = 768, intermediate: int = 3072, lanes: int = 2) -> int:
    """Device bytes one token row of a batch costs a handle (b2_api.cu: alloc_lane): ids + mask, the bf16
    arrays x / ctx / y [H], qkv [3H], ffn [I], two (mean, rstd) pairs and two sets of H/64 partial sums,
    times the number of batch lanes. 1,024 x 128 rows of BERT-base: about 4.1 GB."""
    slots = (hidden + 63) // 64
    return lanes * (5 + 2 * (3 * hidden + 3 * hidden + intermediate) + 2 * 8 + 2 * 8 * slots)


def classify_limits(free_bytes: int, weight_bytes: int = 700 << 20, hidden: int = 768, intermediate: int = 3072) -> dict:
    """The largest `batch` x `max_len` a new handle fits into `free_bytes` of HBM (weights and a 1 GiB margin
    set aside), as rows and as the batch it allows at the two padded lengths of BASELINE.json; `batch` is
    capped at the library's 4,096 items per encoder pass."""
    per_row = activation_bytes_per_row(hidden, intermediate)
    rows = max(0, (int(free_bytes) - int(weight_bytes) - (1 << 30)) // per_row)
    return {"bytes_per_token_row": per_row, "max_token_rows": int(rows),
            "max_batch_at_128": int(min(4096, rows // 128)), "max_batch_at_512": int(min(4096, rows // 512))}


def _local_device() -> int:
    """One process per GPU under torchrun: LOCAL_RANK picks the device; otherwise $B2_DEVICE or 0."""
    for var in ("B2_DEVICE", "LOCAL_RANK"):
        v = os.environ.get(var)
        if v is not None and v.strip().lstrip("-").isdigit():
            return int(v)
    return 0


def job_devices() -> List[int]:
    """The devices one op call spreads a job over.

    The agent runs tasks inline in one process (app.py:286), so by default that process drives every
    GPU of the box through an EngineGroup. $B2_DEVICES narrows it ("0,1,2,3", or "all"); a process
    that was started one-per-GPU (torchrun sets LOCAL_RANK, or $B2_DEVICE is given) keeps to its own
    device — there the slices and the all-reduce belong to torch.distributed (ops/_shard.py)."""
    if any(os.environ.get(v, "").strip().lstrip("-").isdigit() for v in ("B2_DEVICE", "LOCAL_RANK")):
        return [_local_device()]
    n = device_count()
    spec = (os.environ.get("B2_DEVICES") or "all").strip().lower()
    if spec in ("all", "*", ""):
        return list(range(n))
    devs = [int(x) for x in spec.replace(" ", "").split(",") if x != ""]
    bad = [d for d in devs if d < 0 or d >= n]
    if bad or not devs:
        raise ValueError(f"B2_DEVICES={spec!r} names devices outside [0, {n})")
    return devs


def get_cuda_handle(model_path: str, vocab_path: Optional[str] = None, max_len: int = 128,
                    max_batch: int = 1024) -> CUDAHandle:
    """Create/reuse the engine for `model_path` (ops/_tpu_runtime.py:34-63).

    Imports (dlopens) the native library only when needed, so the agent can start without it."""
    global _ENGINE, _INPUT_DETAILS, _OUTPUT_DETAILS, _MODEL_PATH, _CACHE_KEY

    vocab = get_vocab_path(model_path, vocab_path)
    key = (model_path, vocab, int(max_len), int(max_batch))
    with _LOCK:
        # Reuse if the same model is already loaded
        if _ENGINE is not None and _CACHE_KEY is not None and _CACHE_KEY[:2] == key[:2] \
                and _CACHE_KEY[2] >= key[2] and _CACHE_KEY[3] >= key[3]:
            return CUDAHandle(_ENGINE, _INPUT_DETAILS, _OUTPUT_DETAILS, _MODEL_PATH)

        lib = load_library()
        if lib.b2_device_count() <= 0:
            raise RuntimeError("No CUDA device detected (b2_device_count() returned 0).")
        if not os.path.exists(model_path):
            raise FileNotFoundError(f"CUDA model not found: {model_path}")
        if not os.path.exists(vocab):
            raise FileNotFoundError(f"CUDA tokenizer vocabulary not found: {vocab}")

        if _ENGINE is not None:
            if _CSV_INDEX is not None and _CSV_INDEX._engine is _ENGINE:
                _drop_csv_index()
            _ENGINE.close()
            _ENGINE = None
        devs = job_devices()
        if len(devs) > 1:
            eng = EngineGroup(model_path, vocab, devices=devs, max_batch=max_batch, max_len=max_len)
        else:
            eng = Engine(model_path, vocab, device=devs[0], max_batch=max_batch, max_len=max_len)
        _ENGINE = eng
        _INPUT_DETAILS = [{"index": 0, "shape": (1, int(eng.info.max_len)), "dtype": np.int32}]
        _OUTPUT_DETAILS = [{"index": 0, "shape": (1, int(eng.info.num_labels)), "dtype": np.float32}]
        _MODEL_PATH = model_path
        _CACHE_KEY = (model_path, vocab, int(eng.info.max_len), int(eng.info.max_batch))
        return CUDAHandle(_ENGINE, _INPUT_DETAILS, _OUTPUT_DETAILS, _MODEL_PATH)


# A tokenizer-only engine for map_tokenize (no model needed): cached per vocabulary path.
_TOK_ENGINE = None
_TOK_KEY = None


def get_tokenizer_engine(vocab_path: Optional[str] = None, max_len: int = 512) -> Engine:
    global _TOK_ENGINE, _TOK_KEY
    with _LOCK:
        if _ENGINE is not None and vocab_path in (None, _ENGINE.vocab_path) and _ENGINE.info.max_len >= max_len:
            return _ENGINE
        key = (vocab_path, int(max_len))
        if _TOK_ENGINE is not None and _TOK_KEY[0] == vocab_path and _TOK_KEY[1] >= max_len:
            return _TOK_ENGINE
        lib = load_library()
        if lib.b2_device_count() <= 0:
            raise RuntimeError("No CUDA device detected (b2_device_count() returned 0).")
        if vocab_path is not None and not os.path.exists(vocab_path):
            raise FileNotFoundError(f"CUDA tokenizer vocabulary not found: {vocab_path}")
        if _TOK_ENGINE is not None:
            if _CSV_INDEX is not None and _CSV_INDEX._engine is _TOK_ENGINE:
                _drop_csv_index()
            _TOK_ENGINE.close()
        _TOK_ENGINE = Engine(None, vocab_path, device=_local_device(), max_batch=1, max_len=max_len)
        _TOK_KEY = key
        return _TOK_ENGINE


# The last CSV file indexed, kept resident so that successive shard tasks over one file pay for the
# copy and the scan once (the reference re-parses from the top of the file for every shard).
_CSV_INDEX = None
_CSV_KEY = None
_CSV_MAP = None


def get_csv_index(path: str, engine: Engine):
    """(CsvIndex, read-only mmap of the file) for `path` on `engine`; single-slot cache keyed on the
    file's identity (real path, size, mtime) and the engine."""
    global _CSV_INDEX, _CSV_KEY, _CSV_MAP
    import mmap

    st = os.stat(path)
    key = (os.path.realpath(path), st.st_size, st.st_mtime_ns, id(engine), engine.handle.value)
    with _LOCK:
        if _CSV_INDEX is not None and _CSV_KEY == key:
            return _CSV_INDEX, _CSV_MAP
        _drop_csv_index()
        with open(path, "rb") as f:
            data = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ) if st.st_size > 0 else b""
        _CSV_INDEX = CsvIndex(engine, data)
        _CSV_MAP = data
        _CSV_KEY = key
        return _CSV_INDEX, _CSV_MAP


def _drop_csv_index() -> None:
    global _CSV_INDEX, _CSV_KEY, _CSV_MAP
    if _CSV_INDEX is not None:
        _CSV_INDEX.close()
    if _CSV_MAP is not None and hasattr(_CSV_MAP, "close"):
        try:
            _CSV_MAP.close()
        except BufferError:
            pass
    _CSV_INDEX = _CSV_KEY = _CSV_MAP = None


def reset_cache() -> None:
    """Drop cached engines (tests)."""
    global _ENGINE, _INPUT_DETAILS, _OUTPUT_DETAILS, _MODEL_PATH, _CACHE_KEY, _TOK_ENGINE, _TOK_KEY
    with _LOCK:
        _drop_csv_index()
        for e in (_ENGINE, _TOK_ENGINE):
            if e is not None:
                e.close()
        _ENGINE = _INPUT_DETAILS = _OUTPUT_DETAILS = _MODEL_PATH = _CACHE_KEY = None
        _TOK_ENGINE = _TOK_KEY = None
