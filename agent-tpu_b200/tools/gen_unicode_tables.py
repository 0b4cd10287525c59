#!/usr/bin/env python3
"""Generate data/bert_unicode.bin — the per-code-point tables the GPU tokenizer uses.

The tables are obtained by PROBING the pinned third-party tokenizer (tokenizers 0.22.2,
`BertNormalizer(clean_text, handle_chinese_chars, strip_accents, lowercase)` and
`BertPreTokenizer`) one code point at a time, so that the CUDA kernel reproduces that library's
Unicode behaviour (its NFD data, its category tables, its lower-casing) without re-deriving it from
a different Unicode version. The reference itself has no tokenizer; this is the
tokenizer the north star asks for.

Per input code point the normaliser emits 0..3 output code points; each output code point is
  class 0  word character
  class 1  whitespace (separates words, emits nothing)
  class 2  isolated   (punctuation or CJK ideograph: always a word of its own)

File layout (little endian):
  magic  'B2UT' u32, version u32 = 1, n_pages u32, n_unique_pages u32, pool_len u32, pad u32
  page_index  u16[n_pages]            (code point >> 7 -> unique page number), padded to 4 bytes
  pages       u32[n_unique_pages*128] entry: bits[0,21) delta = (out - cp) mod 2^21 when n_out == 1,
                                              or pool offset when n_out >= 2;
                                              bits[21,23) class of the single output; bits[23,26) n_out
  pool        u32[pool_len]           entry: bits[0,21) output code point, bits[21,23) class

Limit stated in DESIGN.md: canonical re-ordering of several combining marks that survive the
accent strip (general category Mc/Me with non-zero combining class) is context dependent and is
not modelled; no such sequence occurs in the synthetic corpus.
"""
import struct
import sys
from pathlib import Path

import numpy as np
from tokenizers.normalizers import BertNormalizer
from tokenizers.pre_tokenizers import BertPreTokenizer

PAGE = 128
N_CP = 0x110000


def probe():
    norm = BertNormalizer(clean_text=True, handle_chinese_chars=True, strip_accents=True, lowercase=True)
    pre = BertPreTokenizer()
    cls_cache = {}

    def out_class(ch):
        c = cls_cache.get(ch)
        if c is None:
            parts = [p for p, _ in pre.pre_tokenize_str("a" + ch + "a")]
            if parts == ["a", "a"]:
                c = 1
            elif parts == ["a", ch, "a"]:
                c = 2
            elif parts == ["a" + ch + "a"]:
                c = 0
            else:
                raise RuntimeError(f"unexpected split for U+{ord(ch):04X}: {parts}")
            cls_cache[ch] = c
        return c

    outs = []
    for cp in range(N_CP):
        if 0xD800 <= cp <= 0xDFFF:
            outs.append([])  # surrogates cannot appear in valid UTF-8
            continue
        o = norm.normalize_str(chr(cp))
        if len(o) == 3 and o[0] == " " and o[2] == " " and o[1] != " ":
            # handle_chinese_chars pads the ideograph with spaces: same as isolating it
            outs.append([(ord(o[1]), 2)])
            continue
        outs.append([(ord(ch), out_class(ch)) for ch in o])
    return outs


def build(outs):
    entries = np.zeros(N_CP, dtype=np.uint32)
    pool = []
    pool_index = {}
    for cp, o in enumerate(outs):
        n = len(o)
        if n > 7:
            raise RuntimeError("expansion too long")
        if n == 0:
            e = 0
        elif n == 1:
            out, c = o[0]
            e = ((out - cp) & 0x1FFFFF) | (c << 21) | (1 << 23)
        else:
            key = tuple(o)
            off = pool_index.get(key)
            if off is None:
                off = len(pool)
                pool_index[key] = off
                pool.extend((out | (c << 21)) for out, c in o)
            e = off | (n << 23)
        entries[cp] = e
    pages = entries.reshape(-1, PAGE)
    uniq = {}
    page_index = np.zeros(pages.shape[0], dtype=np.uint16)
    uniq_pages = []
    for i, pg in enumerate(pages):
        k = pg.tobytes()
        j = uniq.get(k)
        if j is None:
            j = len(uniq_pages)
            uniq[k] = j
            uniq_pages.append(pg)
        page_index[i] = j
    return page_index, np.stack(uniq_pages), np.array(pool, dtype=np.uint32)


def main():
    out_path = Path(sys.argv[1]) if len(sys.argv) > 1 else Path(__file__).resolve().parents[1] / "data" / "bert_unicode.bin"
    outs = probe()
    page_index, pages, pool = build(outs)
    pi = page_index.tobytes()
    if len(pi) % 4:
        pi += b"\0" * (4 - len(pi) % 4)
    blob = struct.pack("<4sIIIII", b"B2UT", 1, len(page_index), pages.shape[0], len(pool), 0)
    blob += pi + pages.astype("<u4").tobytes() + pool.astype("<u4").tobytes()
    out_path.parent.mkdir(parents=True, exist_ok=True)
    out_path.write_bytes(blob)
    mx = max(len(o) for o in outs)
    print(f"wrote {out_path}: {len(blob)} bytes, {pages.shape[0]} unique pages, pool {len(pool)}, max expansion {mx}")


if __name__ == "__main__":
    main()
