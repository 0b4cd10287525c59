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


def test_noalloc_leaves_device_memory_untouched_and_checks_offsets(vocab_path, sentences):
    """What the name promises: once the first call on the GPU has looked up the block's address, a
    call changes the amount of free device memory by nothing — while b2_tokenize on a job larger than
    any before it has to grow its staging buffers. And because the kernel indexes the shared block
    with the caller's offsets, offsets that run backwards or past the text are refused up front."""
    from ops import _cuda_runtime as rt
    eng = rt.Engine(None, vocab_path, max_batch=1, max_len=128)
    try:
        rc, _, _, err = _noalloc(eng, rt, sentences[:8], 128)            # warm: address lookup done
        assert rc == 0, err
        free_before, _ = rt.device_memory(eng.device)
        for texts in (sentences[:3000], sentences[:100], sentences[:3000]):
            rc, ids, _, err = _noalloc(eng, rt, texts, 128)
            assert rc == 0, err
            assert rt.device_memory(eng.device)[0] == free_before
        eng.tokenize(sentences[:3000], 128)                              # the allocating path, for contrast
        assert rt.device_memory(eng.device)[0] < free_before
        # malformed offsets
        lib = rt.load_library()
        blob, offsets = rt.pack_items(sentences[:4])
        out = np.zeros((4, 128), dtype=np.int32)
        for bad in (offsets[[0, 2, 1, 3, 4]], np.array([0, 5, 10, 10 ** 9, offsets[4]], dtype=np.int64),
                    offsets + 1):
            bad = np.ascontiguousarray(bad, dtype=np.int64)
            rc = lib.b2_tokenize_noalloc(eng.handle, blob.ctypes.data_as(C.c_void_p), bad.ctypes.data_as(C.c_void_p),
                                         4, 128, out.ctypes.data_as(C.c_void_p), None)
            assert rc == 1, (rc, lib.b2_last_error())
        # sizes whose product would wrap a 64-bit sum are refused, not wrapped
        rc = lib.b2_tokenize_noalloc(eng.handle, blob.ctypes.data_as(C.c_void_p), offsets.ctypes.data_as(C.c_void_p),
                                     4, 2 ** 31 - 1, out.ctypes.data_as(C.c_void_p), None)
        assert rc == 5
        rc, ids, _, err = _noalloc(eng, rt, sentences[:4], 128)          # still healthy
        assert rc == 0 and (ids == eng.tokenize(sentences[:4], 128)[0]).all()
    finally:
        eng.close()
