"""Deterministic synthetic inputs for map_classify / map_tokenize: vocabulary and sentences.

There is no network and no word list on the box, and the reference ships neither a vocabulary nor
a model, so the benchmark and the tests use:

  * a 30,522-entry WordPiece vocabulary with BERT's special-token ids ([PAD]=0, [UNK]=100,
    [CLS]=101, [SEP]=102, [MASK]=103), printable ASCII singles and their "##" forms, every
    syllable of the generator as a word-start and as a "##" piece, and the most frequent words of a
    50,000-word synthetic lexicon as whole-word entries;
  * sentences drawn Zipf-like from that lexicon: frequent words are whole-word vocabulary hits,
    rarer ones split into 2-4 pieces, and ~4% carry a letter the vocabulary lacks and become [UNK].

Everything is a pure function of the seed (default 1234).
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import numpy as np

VOCAB_SIZE = 30522
LEXICON_SIZE = 50000
SEED = 1234

_ONSETS = list("bcdfghjklmnprstvwz") + ["ch", "sh", "th", "br", "st", "tr", "pl", "gr"]
_NUCLEI = list("aeiou") + ["ai", "ea", "ou", "io"]
_CODAS = ["", "n", "r", "s", "t", "l", "m", "k"]
_UNK_LETTERS = ["ʒ", "ћ", "ք"]  # ʒ ћ ք : lower-case letters absent from the vocabulary


def syllables() -> List[str]:
    return [o + n + c for o in _ONSETS for n in _NUCLEI for c in _CODAS]


def build_lexicon(seed: int = SEED, size: int = LEXICON_SIZE) -> Tuple[List[str], List[int]]:
    """Return (words, syllable_counts); words are unique and ordered by Zipf rank."""
    rng = np.random.default_rng(seed)
    syl = syllables()
    words: List[str] = []
    counts: List[int] = []
    seen = set()
    while len(words) < size:
        n = int(rng.choice([1, 2, 3, 4], p=[0.10, 0.45, 0.33, 0.12]))
        w = "".join(syl[int(i)] for i in rng.integers(0, len(syl), size=n))
        if w in seen:
            continue
        seen.add(w)
        words.append(w)
        counts.append(n)
    return words, counts


def build_vocab(seed: int = SEED, size: int = VOCAB_SIZE) -> List[str]:
    """The synthetic vocab.txt as a list (index = token id)."""
    vocab: List[str] = ["[PAD]"] + [f"[unused{i}]" for i in range(99)]
    vocab += ["[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    vocab += [f"[unused{i}]" for i in range(99, 99 + 999 - len(vocab))]
    assert len(vocab) == 999
    present = set(vocab)

    def add(tok: str) -> None:
        if tok not in present and len(vocab) < size:
            present.add(tok)
            vocab.append(tok)

    for c in range(33, 127):
        add(chr(c))
    for c in "abcdefghijklmnopqrstuvwxyz0123456789":
        add("##" + c)
    for ch in "中文日本語汉字":  # a few CJK ideographs as whole tokens
        add(ch)
    for s in syllables():
        add(s)
        add("##" + s)
    words, _ = build_lexicon(seed)
    for w in words:
        if len(vocab) >= size:
            break
        add(w)
    assert len(vocab) == size, len(vocab)
    return vocab


def write_vocab(path: str, seed: int = SEED) -> str:
    vocab = build_vocab(seed)
    tmp = path + ".tmp"
    with open(tmp, "w", encoding="utf-8", newline="\n") as f:
        f.write("\n".join(vocab) + "\n")
    os.replace(tmp, path)
    return path


class SentenceGenerator:
    """Seeded generator of English-like sentences with a controlled WordPiece length."""

    def __init__(self, seed: int = SEED, zipf_a: float = 1.1):
        self.words, self.syl_counts = build_lexicon(seed)
        vocab = set(build_vocab(seed))
        self.in_vocab = np.array([w in vocab for w in self.words])
        ranks = np.arange(1, len(self.words) + 1, dtype=np.float64)
        p = ranks ** (-zipf_a)
        # flatten the head a little so that rarer (multi-piece) words show up in every sentence
        p = np.minimum(p, p[200])
        self.p = p / p.sum()
        self.cdf = np.cumsum(self.p)
        # upper bound of pieces per word: 1 if in the vocabulary else its syllable count (+1 slack)
        self.max_pieces = np.where(self.in_vocab, 1, np.array(self.syl_counts) + 1)

    def sentences(self, n: int, min_tokens: int = 16, max_tokens: int = 126, seed: int = SEED) -> List[str]:
        rng = np.random.default_rng(seed + 7919)
        out: List[str] = []
        targets = rng.integers(min_tokens, max_tokens + 1, size=n)
        for t in targets:
            budget = int(t)
            parts: List[str] = []
            used = 0
            first = True
            while True:
                r = rng.random(3)
                wi = int(np.searchsorted(self.cdf, r[0]))
                wi = min(wi, len(self.words) - 1)
                w = self.words[wi]
                cost = int(self.max_pieces[wi])
                if r[1] < 0.04:  # a letter outside the vocabulary -> the whole word becomes [UNK]
                    k = int(r[2] * len(_UNK_LETTERS)) % len(_UNK_LETTERS)
                    w = w + _UNK_LETTERS[k]
                    cost = 1
                punct = ""
                if r[2] < 0.10:
                    punct = ","
                elif r[2] < 0.16:
                    punct = "."
                total = cost + (1 if punct else 0)
                if used + total > budget - 1:  # keep one token for the final full stop
                    break
                if first:
                    w = w[0].upper() + w[1:]
                    first = False
                parts.append(w + punct)
                if punct == ".":
                    first = True
                used += total
            text = " ".join(parts)
            if not text.endswith("."):
                text += "."
            out.append(text)
        return out


def pack_items(items: Sequence[str]) -> Tuple[bytes, np.ndarray]:
    """Concatenate items as UTF-8 and return (bytes, int64 offsets[n+1]) — the C ABI's text layout."""
    enc = [s.encode("utf-8", "surrogatepass") for s in items]
    offsets = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum([len(b) for b in enc], out=offsets[1:])
    return b"".join(enc), offsets
