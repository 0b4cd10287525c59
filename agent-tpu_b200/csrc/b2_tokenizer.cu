// b2_tokenizer.cu — BERT tokenize-and-pad on the GPU: UTF-8 decode, BertNormalizer (clean text,
// CJK isolation, accent strip, lower-case), BertPreTokenizer (whitespace / punctuation split),
// greedy longest-match WordPiece, [CLS] … [SEP], truncation and padding — one warp per line.
//
// The reference has no tokenizer of this kind: ops/map_tokenize.py:6-9 cuts text into fixed-width
// chunks (that behaviour is the chunk kernel at the bottom of this file). This kernel is the
// "WordPiece tokenize-and-pad kernel" of BASELINE.json's north star and is held bit-exact to
// tokenizers 0.22.2 `BertWordPieceTokenizer` (tests/test_gpu_tokenizer.py).
//
// Layout of the work: a warp streams its line through a window of <= 320 normalised characters in
// shared memory. Per window: (1) 32 bytes per step are decoded and normalised in parallel (table
// lookups, warp prefix sum for the output positions); (2) word starts / ends are found with
// ballots and compacted; (3) one lane per word runs the greedy longest-match against a hash table
// of the vocabulary (O(1) substring hashes from per-word prefix hashes, exact code-point compare on
// a hit) and records each piece id at the piece's first character; (4) the piece ids are compacted
// in order into the output row. A word cut by the window edge is carried into the next window.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "b2_tokenizer.h"

namespace b2 {

namespace {

constexpr int NB = 320;            // window capacity in normalised characters
constexpr int MAX_EXP = 3;         // max outputs of one input code point (gen_unicode_tables.py)
constexpr int MAX_WORD = 100;      // WordPiece max_input_chars_per_word
constexpr int TOK_WARPS = 4;
constexpr uint64_t HASH_P = 0x100000001B3ull;
constexpr uint64_t CONT_SALT = 0x9E3779B97F4A7C15ull;

enum { CLS_WORD = 0, CLS_SPACE = 1, CLS_ISO = 2, CLS_SPECIAL = 3, CLS_SKIP = 4 };

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}
// key of a vocabulary piece: polynomial hash of its code points, length and continuation flag.
__host__ __device__ __forceinline__ uint64_t piece_key(uint64_t poly, int len, bool cont) {
  uint64_t k = mix64(poly ^ (cont ? CONT_SALT : 0ull) ^ (static_cast<uint64_t>(len) << 56));
  return k ? k : 1ull;  // 0 marks an empty slot
}

struct WarpBuf {
  uint32_t cps[NB];       // normalised code point (or token id for CLS_SPECIAL)
  uint64_t ph[2 * NB + 1];  // per-word prefix hashes; word number w uses slots [start+w, end+w]
  int32_t pid[NB];        // piece id at the piece's first character, -1 elsewhere
  uint16_t wstart[NB];
  uint16_t wend[NB];      // inclusive last character of the word
  uint8_t cls[NB];
};

__device__ __forceinline__ int lookup_piece(const TokenizerDev& T, uint64_t key, const uint32_t* cps,
                                            int len) {
  uint32_t slot = static_cast<uint32_t>(key >> (64 - T.table_bits));
  const uint32_t mask = (1u << T.table_bits) - 1u;
  for (;;) {
    const VocabSlot s = T.table[slot];
    if (s.key == 0ull) return -1;
    if (s.key == key && static_cast<int>(s.off_len & 0xFFu) == len) {
      const uint32_t* v = T.vocab_cps + (s.off_len >> 8);
      bool same = true;
      for (int i = 0; i < len; ++i) same &= (v[i] == cps[i]);
      if (same) return static_cast<int>(s.id);
    }
    slot = (slot + 1u) & mask;
  }
}

__global__ void __launch_bounds__(TOK_WARPS * 32)
wordpiece_kernel(TokenizerDev T, const uint8_t* __restrict__ text, const int64_t* __restrict__ offsets,
                 int n_lines, int max_len, int32_t* __restrict__ ids, uint8_t* __restrict__ mask,
                 int32_t* __restrict__ lens) {
  __shared__ WarpBuf bufs[TOK_WARPS];
  __shared__ uint64_t pows[MAX_WORD + 1];
  for (int i = threadIdx.x; i <= MAX_WORD; i += blockDim.x) pows[i] = T.pows[i];
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  const unsigned lt_mask = (1u << lane) - 1u;
  WarpBuf& W = bufs[warp];
  const int cap = max_len - 2;

  for (int line = blockIdx.x * TOK_WARPS + warp; line < n_lines; line += gridDim.x * TOK_WARPS) {
    const int64_t beg = offsets[line], end = offsets[line + 1];
    int32_t* row = ids + static_cast<size_t>(line) * max_len;
    int64_t in_pos = beg;
    int64_t skip_until = beg;  // raw bytes below this belong to a matched special token
    int fill = 0;              // characters in the window (the first `fill` may be a carried word)
    int ntok = 0;              // content tokens written so far
    bool skip_cont = false;    // the window starts inside a >100-character word already emitted

    while ((in_pos < end || fill > 0) && ntok < cap) {
      // ---- (1) decode + normalise until the window is nearly full or the line ends ----
      while (in_pos < end && fill + 32 * MAX_EXP <= NB) {
        const int64_t i = in_pos + lane;
        uint32_t b0 = (i < end) ? text[i] : 0x80u;  // 0x80: a continuation byte, emits nothing
        int n_out = 0;
        uint32_t out[MAX_EXP];
        uint32_t oc[MAX_EXP];
        // literal special tokens in the raw text ("[CLS]", "[SEP]", ...): matched before normalising
        int mlen = 0, mid = 0;
        if (b0 == '[') {
          for (int s = 0; s < T.n_special; ++s) {
            const int L = T.special_len[s];
            bool ok = (i + L <= end);
            for (int k = 1; ok && k < L; ++k) ok = (text[i + k] == T.special_text[s][k]);
            if (ok) { mlen = L; mid = T.special_id[s]; break; }
          }
        }
        // covered: this byte lies inside a match that started earlier (matches are <= 8 bytes)
        bool covered = i < skip_until;
        {
#pragma unroll
          for (int j = 1; j < 8; ++j) {
            const int ml = __shfl_up_sync(FULL, mlen, j);
            if (lane >= j && ml > j) covered = true;
          }
        }
        if (mlen > 0 && !covered) {
          n_out = 1; out[0] = static_cast<uint32_t>(mid); oc[0] = CLS_SPECIAL;
        } else if (!covered && (b0 & 0xC0u) != 0x80u) {
          uint32_t cp;
          if (b0 < 0x80u) {
            cp = b0;
          } else {
            const int need = b0 >= 0xF0u ? 3 : (b0 >= 0xE0u ? 2 : 1);
            cp = b0 & (0x3Fu >> need);
            bool ok = (b0 >= 0xC0u) && (b0 < 0xF8u) && (i + need < end);
            for (int k = 1; ok && k <= need; ++k) {
              const uint32_t bk = text[i + k];
              ok = (bk & 0xC0u) == 0x80u;
              cp = (cp << 6) | (bk & 0x3Fu);
            }
            if (!ok || cp >= 0x110000u) cp = 0xFFFDu;
          }
          const uint32_t e = T.pages[static_cast<uint32_t>(T.page_index[cp >> 7]) * 128u + (cp & 127u)];
          n_out = static_cast<int>((e >> 23) & 7u);
          if (n_out == 1) {
            out[0] = (cp + (e & 0x1FFFFFu)) & 0x1FFFFFu;
            oc[0] = (e >> 21) & 3u;
          } else if (n_out > 1) {
            const uint32_t* pl = T.pool + (e & 0x1FFFFFu);
            for (int k = 0; k < n_out && k < MAX_EXP; ++k) {
              out[k] = pl[k] & 0x1FFFFFu;
              oc[k] = (pl[k] >> 21) & 3u;
            }
          }
        }
        // the furthest byte any match in this step covers
        {
          long long reach = (mlen > 0 && !covered) ? i + mlen : 0;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const long long r2 = __shfl_xor_sync(FULL, reach, o);
            reach = r2 > reach ? r2 : reach;
          }
          if (reach > skip_until) skip_until = reach;
        }
        // exclusive prefix sum of n_out over the warp
        int incl = n_out;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(FULL, incl, o);
          if (lane >= o) incl += t;
        }
        const int total = __shfl_sync(FULL, incl, 31);
        int at = fill + incl - n_out;
        for (int k = 0; k < n_out; ++k) { W.cps[at + k] = out[k]; W.cls[at + k] = static_cast<uint8_t>(oc[k]); }
        fill += total;
        in_pos += 32;
      }
      __syncwarp();
      const bool at_end = in_pos >= end;

      // ---- the head of the window may continue an over-long word that was already emitted ----
      if (skip_cont) {
        // first position that is not a plain word character
        int first_non = fill;
        for (int base = 0; base < fill; base += 32) {
          const int i = base + lane;
          const bool nonword = (i < fill) && (W.cls[i] != CLS_WORD);
          const unsigned bal = __ballot_sync(FULL, nonword);
          if (bal) { first_non = base + __ffs(bal) - 1; break; }
        }
        for (int i = lane; i < first_non; i += 32) W.cls[i] = CLS_SKIP;
        __syncwarp();
        if (first_non < fill) skip_cont = false;
      }

      // ---- (2) word starts / ends, compacted in order ----
      int nw = 0, nwe = 0;
      for (int base = 0; base < fill; base += 32) {
        const int i = base + lane;
        bool st = false, la = false;
        if (i < fill) {
          const int c = W.cls[i];
          if (c == CLS_WORD) {
            st = (i == 0) || (W.cls[i - 1] != CLS_WORD);
            la = (i + 1 == fill) || (W.cls[i + 1] != CLS_WORD);
          } else if (c == CLS_ISO || c == CLS_SPECIAL) {
            st = la = true;
          }
        }
        const unsigned bs = __ballot_sync(FULL, st), bl = __ballot_sync(FULL, la);
        if (st) W.wstart[nw + __popc(bs & lt_mask)] = static_cast<uint16_t>(i);
        if (la) W.wend[nwe + __popc(bl & lt_mask)] = static_cast<uint16_t>(i);
        nw += __popc(bs);
        nwe += __popc(bl);
      }
      __syncwarp();
      // a word touching the window edge is incomplete unless the line is finished
      int region_end = fill;
      if (!at_end && fill > 0 && W.cls[fill - 1] == CLS_WORD && nw > 0) {
        const int s = W.wstart[nw - 1];
        if (fill - s <= MAX_WORD) {
          nw -= 1;
          region_end = s;
        } else {
          skip_cont = true;  // > 100 characters: [UNK] now, swallow the rest of the word later
        }
      }

      // ---- (3) WordPiece, one lane per word ----
      for (int i = lane; i < region_end; i += 32) W.pid[i] = -1;
      __syncwarp();
      for (int w0 = 0; w0 < nw; w0 += 32) {
        const int w = w0 + lane;
        if (w < nw) {
          const int s = W.wstart[w], e = W.wend[w] + 1, len = e - s;
          if (W.cls[s] == CLS_SPECIAL) {
            W.pid[s] = static_cast<int32_t>(W.cps[s]);
          } else if (len > MAX_WORD) {
            W.pid[s] = T.unk_id;
          } else {
            // prefix hashes of this word; shifting by the word number keeps adjacent words'
            // slot ranges [s+w, e+w] disjoint (each word needs len+1 slots)
            uint64_t* ph = W.ph + w;
            uint64_t h = 0;
            ph[s] = 0;
            for (int k = s; k < e; ++k) { h = h * HASH_P + (W.cps[k] + 1ull); ph[k + 1] = h; }
            int start = s;
            bool bad = false;
            while (start < e) {
              int stop = e, found = -1;
              while (start < stop) {
                const int L = stop - start;
                const uint64_t poly = ph[stop] - ph[start] * pows[L];
                found = lookup_piece(T, piece_key(poly, L, start > s), &W.cps[start], L);
                if (found >= 0) break;
                --stop;
              }
              if (found < 0) { bad = true; break; }
              W.pid[start] = found;
              start = stop;
            }
            if (bad) {
              for (int k = s + 1; k < e; ++k) W.pid[k] = -1;
              W.pid[s] = T.unk_id;
            }
          }
        }
      }
      __syncwarp();

      // ---- (4) compact piece ids into the output row ----
      for (int base = 0; base < region_end && ntok < cap; base += 32) {
        const int i = base + lane;
        const int v = (i < region_end) ? W.pid[i] : -1;
        const unsigned bal = __ballot_sync(FULL, v >= 0);
        const int at = ntok + __popc(bal & lt_mask);
        if (v >= 0 && at < cap) row[1 + at] = v;
        ntok += __popc(bal);
      }
      if (ntok > cap) ntok = cap;

      // ---- carry the cut word to the front of the window ----
      const int carry = fill - region_end;
      if (carry > 0 && region_end > 0) {
        // forward copy in 32-wide steps; source is always ahead of destination
        for (int base = 0; base < carry; base += 32) {
          const int i = base + lane;
          uint32_t c = 0; uint8_t k = 0;
          if (i < carry) { c = W.cps[region_end + i]; k = W.cls[region_end + i]; }
          __syncwarp();
          if (i < carry) { W.cps[i] = c; W.cls[i] = k; }
          __syncwarp();
        }
      }
      fill = carry;
      if (at_end && carry == 0) break;
      __syncwarp();
    }

    // ---- [CLS] ... [SEP], padding, mask ----
    const int total = ntok + 2;
    if (lane == 0) {
      row[0] = T.cls_id;
      row[1 + ntok] = T.sep_id;
      if (lens) lens[line] = total;
    }
    for (int i = total + lane; i < max_len; i += 32) row[i] = T.pad_id;
    if (mask) {
      uint8_t* mrow = mask + static_cast<size_t>(line) * max_len;
      for (int i = lane; i < max_len; i += 32) mrow[i] = i < total ? 1 : 0;
    }
    __syncwarp();
  }
}

// ---- fixed-width chunking in code points (ops/map_tokenize.py:6-9) ------------------------------
// pass 1: code points per item (bytes that are not UTF-8 continuation bytes); one warp per item.
__global__ void count_codepoints_kernel(const uint8_t* __restrict__ text,
                                        const int64_t* __restrict__ offsets, int n,
                                        int64_t* __restrict__ counts) {
  const int item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= n) return;
  const int lane = threadIdx.x & 31;
  const int64_t beg = offsets[item], end = offsets[item + 1];
  long long c = 0;
  for (int64_t i = beg + lane; i < end; i += 32) c += (text[i] & 0xC0u) != 0x80u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) counts[item] = c;
}
// pass 2: byte offset (relative to the start of `text`) at which chunk j of the item starts, for
// j = 0 .. n_chunks-1, written at bounds[chunk_base[item] + j]; one warp per item.
__global__ void chunk_bounds_kernel(const uint8_t* __restrict__ text,
                                    const int64_t* __restrict__ offsets, int n, int64_t chunk_size,
                                    const int64_t* __restrict__ chunk_base,
                                    int64_t* __restrict__ bounds) {
  const int item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= n) return;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  const int64_t beg = offsets[item], end = offsets[item + 1];
  int64_t* out = bounds + chunk_base[item];
  long long seen = 0;  // code points before this 32-byte step
  for (int64_t b = beg; b < end; b += 32) {
    const int64_t i = b + lane;
    const bool lead = (i < end) && ((text[i] & 0xC0u) != 0x80u);
    const unsigned bal = __ballot_sync(0xffffffffu, lead);
    if (lead) {
      const long long idx = seen + __popc(bal & lt_mask);  // index of this code point in the item
      if (idx % chunk_size == 0) out[idx / chunk_size] = i;
    }
    seen += __popc(bal);
  }
}

// ---- host: vocabulary and table loading ---------------------------------------------------------
bool read_file(const char* path, std::vector<uint8_t>* out) {
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out->resize(n > 0 ? n : 0);
  size_t got = n > 0 ? fread(out->data(), 1, n, f) : 0;
  fclose(f);
  return got == static_cast<size_t>(n > 0 ? n : 0);
}

void utf8_decode(const std::string& s, std::vector<uint32_t>* cps) {
  cps->clear();
  size_t i = 0;
  while (i < s.size()) {
    uint32_t b0 = static_cast<uint8_t>(s[i]);
    int need = b0 < 0x80 ? 0 : (b0 >= 0xF0 ? 3 : (b0 >= 0xE0 ? 2 : 1));
    uint32_t cp = need ? (b0 & (0x3Fu >> need)) : b0;
    for (int k = 1; k <= need && i + k < s.size(); ++k) cp = (cp << 6) | (static_cast<uint8_t>(s[i + k]) & 0x3Fu);
    cps->push_back(cp);
    i += need + 1;
  }
}

template <typename T>
int upload(const std::vector<T>& h, T** d) {
  B2_CUDA_OK(cudaMalloc(d, std::max<size_t>(h.size(), 1) * sizeof(T)));
  if (!h.empty()) B2_CUDA_OK(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return B2_OK;
}

}  // namespace

int tokenizer_load(const char* vocab_path, const char* unicode_path, TokenizerDev* T, int* vocab_size) {
  memset(T, 0, sizeof(*T));
  std::vector<uint8_t> raw;
  if (!unicode_path || !read_file(unicode_path, &raw) || raw.size() < 24) {
    set_last_error("tokenizer: cannot read unicode table file '%s'", unicode_path ? unicode_path : "(null)");
    return B2_ERR_IO;
  }
  uint32_t hdr[6];
  memcpy(hdr, raw.data(), 24);
  if (memcmp(raw.data(), "B2UT", 4) != 0 || hdr[1] != 1) {
    set_last_error("tokenizer: '%s' is not a B2UT v1 table", unicode_path);
    return B2_ERR_FORMAT;
  }
  const uint32_t n_pages = hdr[2], n_unique = hdr[3], pool_len = hdr[4];
  size_t pi_bytes = (static_cast<size_t>(n_pages) * 2 + 3) & ~size_t(3);
  size_t need = 24 + pi_bytes + static_cast<size_t>(n_unique) * 128 * 4 + static_cast<size_t>(pool_len) * 4;
  if (n_pages != 0x110000 / 128 || raw.size() < need) {
    set_last_error("tokenizer: unicode table '%s' is truncated or has the wrong page count", unicode_path);
    return B2_ERR_FORMAT;
  }
  std::vector<uint16_t> page_index(n_pages);
  memcpy(page_index.data(), raw.data() + 24, n_pages * 2);
  std::vector<uint32_t> pages(static_cast<size_t>(n_unique) * 128), pool(pool_len);
  memcpy(pages.data(), raw.data() + 24 + pi_bytes, pages.size() * 4);
  memcpy(pool.data(), raw.data() + 24 + pi_bytes + pages.size() * 4, pool.size() * 4);

  std::vector<uint8_t> vraw;
  if (!vocab_path || !read_file(vocab_path, &vraw)) {
    set_last_error("tokenizer: cannot read vocabulary file '%s'", vocab_path ? vocab_path : "(null)");
    return B2_ERR_IO;
  }
  // one token per line, id = line number (BertWordPieceTokenizer's vocab.txt convention)
  std::vector<std::string> tokens;
  {
    std::string cur;
    for (uint8_t c : vraw) {
      if (c == '\n') { if (!cur.empty() && cur.back() == '\r') cur.pop_back(); tokens.push_back(cur); cur.clear(); }
      else cur.push_back(static_cast<char>(c));
    }
    if (!cur.empty()) tokens.push_back(cur);
  }
  if (tokens.empty()) {
    set_last_error("tokenizer: vocabulary '%s' is empty", vocab_path);
    return B2_ERR_FORMAT;
  }
  int bits = 10;
  while ((1u << bits) < tokens.size() * 4) ++bits;
  std::vector<VocabSlot> table(size_t(1) << bits);
  memset(table.data(), 0, table.size() * sizeof(VocabSlot));
  std::vector<uint32_t> vocab_cps;
  std::vector<uint32_t> cps;
  T->unk_id = T->cls_id = T->sep_id = T->pad_id = -1;
  // later duplicates of a token override earlier ones, as in a Python dict built from the file
  std::vector<int> final_id(tokens.size());
  {
    std::vector<std::pair<std::string, int>> order;
    order.reserve(tokens.size());
    for (size_t i = 0; i < tokens.size(); ++i) order.emplace_back(tokens[i], static_cast<int>(i));
    std::stable_sort(order.begin(), order.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    for (size_t i = 0; i < order.size(); ++i) {
      size_t j = i;
      while (j + 1 < order.size() && order[j + 1].first == order[i].first) ++j;
      for (size_t k = i; k <= j; ++k) final_id[order[k].second] = (k == j) ? order[j].second : -1;
      i = j;
    }
  }
  static const char* kSpecial[5] = {"[UNK]", "[SEP]", "[CLS]", "[PAD]", "[MASK]"};
  T->n_special = 0;
  for (size_t id = 0; id < tokens.size(); ++id) {
    if (final_id[id] < 0) continue;
    const std::string& tok = tokens[id];
    if (tok == "[UNK]") T->unk_id = static_cast<int>(id);
    if (tok == "[CLS]") T->cls_id = static_cast<int>(id);
    if (tok == "[SEP]") T->sep_id = static_cast<int>(id);
    if (tok == "[PAD]") T->pad_id = static_cast<int>(id);
    for (int s = 0; s < 5; ++s) {
      if (tok == kSpecial[s]) {
        const int k = T->n_special++;
        T->special_len[k] = static_cast<int>(tok.size());
        memset(T->special_text[k], 0, 8);
        memcpy(T->special_text[k], tok.data(), tok.size());
        T->special_id[k] = static_cast<int>(id);
      }
    }
    bool cont = tok.size() > 2 && tok[0] == '#' && tok[1] == '#';
    utf8_decode(cont ? tok.substr(2) : tok, &cps);
    if (cps.empty() || cps.size() > MAX_WORD) continue;  // can never match a word
    uint64_t poly = 0;
    for (uint32_t c : cps) poly = poly * HASH_P + (c + 1ull);
    const uint64_t key = piece_key(poly, static_cast<int>(cps.size()), cont);
    uint32_t slot = static_cast<uint32_t>(key >> (64 - bits));
    while (table[slot].key != 0) slot = (slot + 1) & ((1u << bits) - 1);
    table[slot].key = key;
    table[slot].id = static_cast<uint32_t>(id);
    table[slot].off_len = (static_cast<uint32_t>(vocab_cps.size()) << 8) | static_cast<uint32_t>(cps.size());
    vocab_cps.insert(vocab_cps.end(), cps.begin(), cps.end());
    if (vocab_cps.size() >= (1u << 24)) {
      set_last_error("tokenizer: vocabulary too large");
      return B2_ERR_LIMIT;
    }
  }
  if (T->unk_id < 0 || T->cls_id < 0 || T->sep_id < 0 || T->pad_id < 0) {
    set_last_error("tokenizer: vocabulary must contain [UNK], [CLS], [SEP] and [PAD]");
    return B2_ERR_FORMAT;
  }
  std::vector<uint64_t> pows(MAX_WORD + 1);
  pows[0] = 1;
  for (int i = 1; i <= MAX_WORD; ++i) pows[i] = pows[i - 1] * HASH_P;

  int rc;
  if ((rc = upload(page_index, const_cast<uint16_t**>(&T->page_index)))) return rc;
  if ((rc = upload(pages, const_cast<uint32_t**>(&T->pages)))) return rc;
  if ((rc = upload(pool, const_cast<uint32_t**>(&T->pool)))) return rc;
  if ((rc = upload(table, const_cast<VocabSlot**>(&T->table)))) return rc;
  if ((rc = upload(vocab_cps, const_cast<uint32_t**>(&T->vocab_cps)))) return rc;
  if ((rc = upload(pows, const_cast<uint64_t**>(&T->pows)))) return rc;
  T->table_bits = bits;
  if (vocab_size) *vocab_size = static_cast<int>(tokens.size());
  return B2_OK;
}

void tokenizer_free(TokenizerDev* T) {
  cudaFree(const_cast<uint16_t*>(T->page_index));
  cudaFree(const_cast<uint32_t*>(T->pages));
  cudaFree(const_cast<uint32_t*>(T->pool));
  cudaFree(const_cast<VocabSlot*>(T->table));
  cudaFree(const_cast<uint32_t*>(T->vocab_cps));
  cudaFree(const_cast<uint64_t*>(T->pows));
  memset(T, 0, sizeof(*T));
}

int tokenizer_run(cudaStream_t stream, const TokenizerDev& T, const uint8_t* text,
                  const int64_t* offsets, int n_lines, int max_len, int32_t* ids, uint8_t* mask,
                  int32_t* lens, int num_sms) {
  if (max_len < 2) {
    set_last_error("tokenize: max_len must be >= 2 (room for [CLS] and [SEP])");
    return B2_ERR_ARG;
  }
  if (n_lines <= 0) return B2_OK;
  int blocks = (n_lines + TOK_WARPS - 1) / TOK_WARPS;
  const int cap = num_sms * 8;  // grid-stride beyond 8 resident CTAs per SM
  if (blocks > cap) blocks = cap;
  wordpiece_kernel<<<blocks, TOK_WARPS * 32, 0, stream>>>(T, text, offsets, n_lines, max_len, ids, mask,
                                                          lens);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int chunk_count(cudaStream_t stream, const uint8_t* text, const int64_t* offsets, int n,
                int64_t* counts) {
  if (n <= 0) return B2_OK;
  count_codepoints_kernel<<<(n + 7) / 8, 256, 0, stream>>>(text, offsets, n, counts);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int chunk_bounds(cudaStream_t stream, const uint8_t* text, const int64_t* offsets, int n,
                 int64_t chunk_size, const int64_t* chunk_base, int64_t* bounds) {
  if (n <= 0) return B2_OK;
  chunk_bounds_kernel<<<(n + 7) / 8, 256, 0, stream>>>(text, offsets, n, chunk_size, chunk_base, bounds);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace b2
