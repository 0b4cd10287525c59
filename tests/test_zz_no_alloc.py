"""b2_tokenize_noalloc: the tokenizer for callers that may not allocate GPU memory while running
(fixed memory budget, or no tolerance for the stalls an allocation causes). Its device staging is
a fixed 8 MiB block inside the library, so this test checks what such a caller relies on:
the answers equal b2_tokenize's bit for bit, a second call is not polluted by what the first left
in the shared block, and a job that is too large is refused with a message instead of overrunning.

Collected last on purpose (file name): a failure here must not hide any other test's result."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 8 * 1024 * 1024


def _noalloc(eng, rt, texts, max_len, want_mask=True):
    lib = rt.load_library()
    lib.b2_tokenize_noalloc.restype = C.c_int
    lib.b2_tokenize_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                        C.c_void_p, C.c_void_p]
    blob, offsets = rt.pack_items(texts)
    n = len(texts)
    ids = np.full((n, max_len), -7, dtype=np.int32)
    mask = np.full((n, max_len), 9, dtype=np.uint8) if want_mask else None
    rc = lib.b2_tokenize_noalloc(eng.handle, blob.ctypes.data_as(C.c_void_p), offsets.ctypes.data_as(C.c_void_p),
                                 n, max_len, ids.ctypes.data_as(C.c_void_p),
                                 mask.ctypes.data_as(C.c_void_p) if want_mask else None)
    return rc, ids, mask, (lib.b2_last_error() or b"").decode()


def test_noalloc_equals_tokenize_on_two_calls_in_a_row(vocab_path, sentences):
    from ops import _cuda_runtime as rt
    eng = rt.Engine(None, vocab_path, max_batch=1, max_len=512)
    try:
        first = sentences[:1500] + ["", "Héllo, wörld! 中文 [SEP] " + "z" * 120]
        second = ["short", "", "ʒ unknown"] + sentences[1500:1600]      # smaller: stale rows would show
        for texts, max_len in ((first, 128), (second, 64), (second[:5], 512)):
            rc, ids, mask, err = _noalloc(eng, rt, texts, max_len)
            assert rc == 0, err
            want_ids, want_mask, _ = eng.tokenize(texts, max_len)
            assert (ids == want_ids).all() and (mask == want_mask).all()
        rc, ids, _, err = _noalloc(eng, rt, second, 128, want_mask=False)   # mask is optional
        assert rc == 0 and (ids == eng.tokenize(second, 128)[0]).all()
        rc, ids, _, _ = _noalloc(eng, rt, [], 128)
        assert rc == 0 and ids.shape == (0, 128)
    finally:
        eng.close()


def test_noalloc_refuses_jobs_beyond_its_fixed_block(vocab_path, sentences):
    from ops import _cuda_runtime as rt
    eng = rt.Engine(None, vocab_path, max_batch=1, max_len=512)
    try:
        big = (sentences * 8)[:14000]                    # 14000 x (5 x 128 + text) > 8 MiB
        rc, _, _, err = _noalloc(eng, rt, big, 128)
        assert rc == 5 and "fixed block" in err and str(SCRATCH_BYTES) in err
        ok = big[:4000]                                  # fits
        rc, ids, mask, err = _noalloc(eng, rt, ok, 128)
        assert rc == 0, err
        assert (ids == eng.tokenize(ok, 128)[0]).all()
    finally:
        eng.close()
