"""The C-ABI shared library: it loads, exports every symbol include/b2_classify.h declares, and
fails loudly (no CPU fallback) when there is no CUDA device. No compute calls without a GPU."""
import ctypes as C
import os
import re

import pytest

from ops import _cuda_runtime as rt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b2_classify.h")


def declared_symbols():
    text = open(HEADER, encoding="utf-8").read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("b2_create", "b2_destroy", "b2_last_error", "b2_tokenize", "b2_classify_text",
                 "b2_classify_text_device", "b2_classify_ids", "b2_chunk_count", "b2_chunk_bounds",
                 "b2_device_count", "b2_get_device_info", "b2_get_model_info", "b2_set_profiling",
                 "b2_get_stream", "b2_tokenize_noalloc", "b2_csv_open", "b2_csv_close", "b2_csv_header",
                 "b2_csv_row_spans", "b2_csv_rows_size", "b2_csv_rows_read", "b2_csv_classify"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(rt.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} is declared in b2_classify.h but not exported"


def test_struct_sizes_match_the_header():
    # b2_config: 2 x i32, 3 pointers, 2 x i32 ; b2_model_info: 12 x i32 + 2 x i64
    assert C.sizeof(rt._Config) == 8 + 3 * C.sizeof(C.c_void_p) + 8
    assert C.sizeof(rt.ModelInfo) == 12 * 4 + 16 + 8
    assert C.sizeof(rt.Metrics) == 8 + 7 * 8 + 8 + 10 * 8 + 10 * 8
    assert C.sizeof(rt.DeviceInfo) == 128 + 8 + 16
    # b2_csv_info: 4 x i32, 5 x i64, 2 x i32, 3 x i64, f64, i64, f64
    assert C.sizeof(rt.CsvInfo) == 16 + 40 + 8 + 24 + 8 + 8 + 8


def _no_cuda():
    return rt.device_count() == 0


@pytest.mark.skipif(not _no_cuda(), reason="this checks the no-device behaviour")
def test_no_device_fails_loudly(tmp_path):
    lib = rt.load_library()
    assert lib.b2_device_count() == 0
    with pytest.raises(rt.B2Error) as e:
        rt.Engine(None, None)
    assert "No CUDA device" in str(e.value)
    with pytest.raises(RuntimeError):
        rt.get_cuda_handle(str(tmp_path / "missing.b2m"))


def test_bad_config_is_rejected():
    lib = rt.load_library()
    h = C.c_void_p()
    assert lib.b2_create(None, C.byref(h)) != 0
    assert b"struct_size" in lib.b2_last_error() or b"NULL" in lib.b2_last_error()
