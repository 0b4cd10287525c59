// b2_csv.cu — CSV record splitter and field extractor for the `read_csv_shard` op and the
// csv -> tokenize -> classify chain (reference: ops/csv_shard.py:9-26, which drives Python's
// csv.DictReader over a file opened with newline="" and encoding="utf-8").
//
// What has to be reproduced is the state machine of CPython's csv reader with its default dialect
// (delimiter ',', quotechar '"', doublequote on, no escapechar, strict off), fed by a line iterator
// that ends a physical line after "\n", "\r" or "\r\n". On raw bytes that is a six-state automaton:
//
//   SR   start of record            ACR  start of record right after a '\r' (a '\n' is swallowed)
//   SF   start of field (after ',') IF   inside an unquoted field
//   IQ   inside a quoted field      QQ   just saw a quote inside a quoted field
//
// A record that starts (first byte that is not a line end, seen in SR/ACR) is a row; line ends seen
// in SR/ACR are the reader's empty records, which DictReader drops. Quotes only open a field at its
// start, a quote in the middle of an unquoted field is data, and after a closing quote any other
// byte continues the field unquoted — none of which a quote-parity scan gets right, hence the
// automaton.
//
// The file is split into 64-byte chunks, one per thread. Pass 1 runs every chunk from all six
// start states at once (a transition vector packed into 18 bits) and composes the vectors per
// block; a single-CTA scan turns the block vectors into the true state at every block boundary.
// Pass 2 re-walks each chunk from its now known state and counts row starts; after a scan of the
// counts, pass 3 writes the byte offset of every row start and row end. Everything is integer
// work bound by HBM reads of the file (3 reads of every byte, 4 + 2 bytes of side data per chunk).
#include "b2_kernels.h"

namespace b2 {
namespace {

constexpr int CSV_CHUNK = 64;   // bytes per thread
constexpr int CSV_BLOCK = 256;  // threads per block: 16 KiB of file per block
constexpr int SCAN_THREADS = 1024;

enum : uint32_t { SR = 0, ACR = 1, SF = 2, IF = 3, IQ = 4, QQ = 5 };
enum : uint32_t { C_OTHER = 0, C_QUOTE = 1, C_COMMA = 2, C_LF = 3, C_CR = 4 };

constexpr uint32_t pack6(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f) {
  return a | (b << 3) | (c << 6) | (d << 9) | (e << 12) | (f << 15);
}
// next state, indexed by the current state (3 bits each), one word per byte class
constexpr uint32_t F_OTHER = pack6(IF, IF, IF, IF, IQ, IF);
constexpr uint32_t F_QUOTE = pack6(IQ, IQ, IQ, IF, QQ, IQ);
constexpr uint32_t F_COMMA = pack6(SF, SF, SF, SF, IQ, SF);
constexpr uint32_t F_LF = pack6(SR, SR, SR, SR, IQ, SR);
constexpr uint32_t F_CR = pack6(ACR, ACR, ACR, ACR, IQ, ACR);
constexpr uint32_t F_IDENT = pack6(0, 1, 2, 3, 4, 5);

__device__ __forceinline__ uint32_t byte_class(uint32_t b) {
  return b == '"' ? C_QUOTE : b == ',' ? C_COMMA : b == '\n' ? C_LF : b == '\r' ? C_CR : C_OTHER;
}
__device__ __forceinline__ uint32_t class_fn(uint32_t cls) {
  return cls == C_OTHER ? F_OTHER
                        : cls == C_QUOTE ? F_QUOTE : cls == C_COMMA ? F_COMMA : cls == C_LF ? F_LF : F_CR;
}
__device__ __forceinline__ uint32_t apply(uint32_t fn, uint32_t state) { return (fn >> (3 * state)) & 7u; }
// (a then b)[s] = b[a[s]]
__device__ __forceinline__ uint32_t compose(uint32_t a, uint32_t b) {
  uint32_t r = 0;
#pragma unroll
  for (int s = 0; s < 6; ++s) r |= apply(b, apply(a, s)) << (3 * s);
  return r;
}
// 0x80 in every byte of w that is one of the four structural bytes
__device__ __forceinline__ uint32_t structural_mask(uint32_t w) {
  return __vcmpeq4(w, 0x22222222u) | __vcmpeq4(w, 0x2c2c2c2cu) | __vcmpeq4(w, 0x0a0a0a0au) |
         __vcmpeq4(w, 0x0d0d0d0du);
}

// Ordered composition over the block: returns in thread 0 the composition of all threads' v.
__device__ uint32_t block_compose_all(uint32_t v, uint32_t* smem /* CSV_BLOCK / 32 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t right = __shfl_down_sync(0xffffffffu, v, off);
    if ((lane & (2 * off - 1)) == 0) v = compose(v, right);
  }
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  uint32_t total = F_IDENT;
  if (threadIdx.x == 0)
    for (int w = 0; w < CSV_BLOCK / 32; ++w) total = compose(total, smem[w]);
  return total;
}
// Ordered exclusive scan over the block: composition of the vectors of all threads before this one.
__device__ uint32_t block_compose_exclusive(uint32_t v, uint32_t* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t left = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl = compose(left, incl);
  }
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  uint32_t before = F_IDENT;
  for (int w = 0; w < warp; ++w) before = compose(before, smem[w]);
  uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = F_IDENT;
  return compose(before, excl);
}
__device__ int64_t block_sum_exclusive(int64_t v, int64_t* smem, int64_t* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  int64_t incl = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int64_t left = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += left;
  }
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  int64_t before = 0, all = 0;
  for (int w = 0; w < nwarp; ++w) {
    if (w < warp) before += smem[w];
    all += smem[w];
  }
  if (total) *total = all;
  __syncthreads();
  return before + incl - v;
}

// Walk the bytes [max(lo, first), min(lo + 64, n)) of one chunk from a known state. `on_start(pos)`
// is called at the first byte of a row, `on_end(pos)` at the line end that closes one.
template <typename OnStart, typename OnEnd>
__device__ __forceinline__ uint32_t walk_chunk(const uint8_t* __restrict__ bytes, int64_t lo, int64_t first,
                                               int64_t n, uint32_t state, OnStart&& on_start,
                                               OnEnd&& on_end) {
  const uint4* src = reinterpret_cast<const uint4*>(bytes + lo);
#pragma unroll 1
  for (int q = 0; q < CSV_CHUNK / 16; ++q) {
    const int64_t qlo = lo + q * 16;
    if (qlo >= n) break;
    const uint4 v = __ldg(src + q);
    const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int64_t wlo = qlo + w * 4;
      const uint32_t word = words[w];
      const bool whole = wlo >= first && wlo + 4 <= n;
      if (whole && state >= IF && structural_mask(word) == 0) {
        state = apply(F_OTHER, state);
        continue;
      }
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int64_t pos = wlo + b;
        if (pos < first || pos >= n) continue;
        const uint32_t cls = byte_class((word >> (8 * b)) & 0xffu);
        if (state <= ACR) {
          if (cls < C_LF) on_start(pos);
        } else if (state != IQ && cls >= C_LF) {
          on_end(pos);
        }
        state = apply(class_fn(cls), state);
      }
    }
  }
  return state;
}

// ---- pass 1: transition vector of every chunk, composed per block ------------------------------
__global__ void __launch_bounds__(CSV_BLOCK)
csv_vectors_kernel(const uint8_t* __restrict__ bytes, int64_t first, int64_t n,
                   uint32_t* __restrict__ chunk_vec, uint32_t* __restrict__ block_vec) {
  __shared__ uint32_t sm[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  uint32_t v = F_IDENT;
  bool settled = false;  // v already ends in F_OTHER (which is idempotent)
  if (lo < n && lo + CSV_CHUNK > first) {
    const uint4* src = reinterpret_cast<const uint4*>(bytes + lo);
#pragma unroll 1
    for (int q = 0; q < CSV_CHUNK / 16; ++q) {
      const int64_t qlo = lo + q * 16;
      if (qlo >= n) break;
      const uint4 ld = __ldg(src + q);
      const uint32_t words[4] = {ld.x, ld.y, ld.z, ld.w};
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const int64_t wlo = qlo + w * 4;
        const uint32_t word = words[w];
        if (wlo >= first && wlo + 4 <= n && structural_mask(word) == 0) {
          if (!settled) v = compose(v, F_OTHER);
          settled = true;
          continue;
        }
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int64_t pos = wlo + b;
          if (pos < first || pos >= n) continue;
          const uint32_t cls = byte_class((word >> (8 * b)) & 0xffu);
          if (cls == C_OTHER) {
            if (!settled) v = compose(v, F_OTHER);
            settled = true;
          } else {
            v = compose(v, class_fn(cls));
            settled = false;
          }
        }
      }
    }
  }
  chunk_vec[chunk] = v;
  const uint32_t total = block_compose_all(v, sm);
  if (threadIdx.x == 0) block_vec[blockIdx.x] = total;
}

// ---- single-CTA scans over the per-block results ------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS)
csv_scan_vectors_kernel(const uint32_t* __restrict__ block_vec, int64_t n_blocks, uint32_t init_state,
                        uint8_t* __restrict__ block_state, int32_t* __restrict__ final_state) {
  __shared__ uint32_t sm[SCAN_THREADS / 32];
  const int64_t per = (n_blocks + SCAN_THREADS - 1) / SCAN_THREADS;
  const int64_t lo = min(n_blocks, int64_t(threadIdx.x) * per), hi = min(n_blocks, lo + per);
  uint32_t agg = F_IDENT;
  for (int64_t b = lo; b < hi; ++b) agg = compose(agg, block_vec[b]);
  // exclusive ordered scan over 1024 threads
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = agg;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t left = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl = compose(left, incl);
  }
  if (lane == 31) sm[warp] = incl;
  __syncthreads();
  uint32_t before = F_IDENT;
  for (int w = 0; w < warp; ++w) before = compose(before, sm[w]);
  uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = F_IDENT;
  uint32_t state = apply(compose(before, excl), init_state);
  for (int64_t b = lo; b < hi; ++b) {
    block_state[b] = static_cast<uint8_t>(state);
    state = apply(block_vec[b], state);
  }
  if (threadIdx.x == SCAN_THREADS - 1) *final_state = static_cast<int32_t>(state);
}

__global__ void __launch_bounds__(SCAN_THREADS)
csv_scan_counts_kernel(const uint32_t* __restrict__ block_count, int64_t n_blocks,
                       int64_t* __restrict__ block_base, int64_t* __restrict__ total) {
  __shared__ int64_t sm[SCAN_THREADS / 32];
  const int64_t per = (n_blocks + SCAN_THREADS - 1) / SCAN_THREADS;
  const int64_t lo = min(n_blocks, int64_t(threadIdx.x) * per), hi = min(n_blocks, lo + per);
  int64_t sum = 0;
  for (int64_t b = lo; b < hi; ++b) sum += block_count[b];
  int64_t all = 0;
  int64_t run = block_sum_exclusive(sum, sm, &all);
  for (int64_t b = lo; b < hi; ++b) {
    block_base[b] = run;
    run += block_count[b];
  }
  if (threadIdx.x == 0) *total = all;
}

// ---- pass 2: true start state of every chunk, row starts per chunk ------------------------------
// ACR differs from SR only in what it means for the row count: a '\n' seen there belongs to the
// line end that already closed the previous record, so it is not an empty record of its own. Neither
// state counts its line ends (DictReader drops empty records), so the class tables cover both.
__global__ void __launch_bounds__(CSV_BLOCK)
csv_count_kernel(const uint8_t* __restrict__ bytes, int64_t first, int64_t n,
                 const uint32_t* __restrict__ chunk_vec, const uint8_t* __restrict__ block_state,
                 uint8_t* __restrict__ chunk_state, uint8_t* __restrict__ chunk_starts,
                 uint32_t* __restrict__ block_starts) {
  __shared__ uint32_t sm[CSV_BLOCK / 32];
  __shared__ int64_t sm64[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  const uint32_t prefix = block_compose_exclusive(chunk_vec[chunk], sm);
  const uint32_t state = apply(prefix, block_state[blockIdx.x]);
  int starts = 0;
  if (lo < n && lo + CSV_CHUNK > first)
    walk_chunk(bytes, lo, first, n, state, [&](int64_t) { ++starts; }, [](int64_t) {});
  chunk_state[chunk] = static_cast<uint8_t>(state);
  chunk_starts[chunk] = static_cast<uint8_t>(starts);
  int64_t all = 0;
  block_sum_exclusive(starts, sm64, &all);
  if (threadIdx.x == 0) block_starts[blockIdx.x] = static_cast<uint32_t>(all);
}

// ---- pass 3: byte offsets of row starts and row ends -------------------------------------------
__global__ void __launch_bounds__(CSV_BLOCK)
csv_emit_kernel(const uint8_t* __restrict__ bytes, int64_t first, int64_t n,
                const uint8_t* __restrict__ chunk_state, const uint8_t* __restrict__ chunk_starts,
                const int64_t* __restrict__ block_base, int64_t* __restrict__ row_start,
                int64_t* __restrict__ row_end) {
  __shared__ int64_t sm64[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  int64_t row = block_base[blockIdx.x] + block_sum_exclusive(chunk_starts[chunk], sm64, nullptr);
  if (lo < n && lo + CSV_CHUNK > first)
    walk_chunk(
        bytes, lo, first, n, chunk_state[chunk], [&](int64_t pos) { row_start[row++] = pos; },
        [&](int64_t pos) { row_end[row - 1] = pos; });
}

// ---- UTF-8 validation (the reference decodes the file strictly: ops/csv_shard.py:17) -------------
// first_bad receives the smallest offset at which a strict decoder would have the evidence of an
// error in hand: the stray or illegal byte itself, the continuation byte that breaks a sequence,
// or n when the data ends inside a sequence.
__device__ __forceinline__ int utf8_len(uint32_t b) {
  if (b < 0x80) return 1;
  if (b < 0xc2) return 0;  // continuation byte or overlong 2-byte lead
  if (b < 0xe0) return 2;
  if (b < 0xf0) return 3;
  if (b < 0xf5) return 4;
  return 0;
}
__global__ void __launch_bounds__(256)
utf8_validate_kernel(const uint8_t* __restrict__ bytes, int64_t n, unsigned long long* first_bad) {
  const int64_t lo = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) * 16;
  if (lo >= n) return;
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(bytes + lo));
  if (((v.x | v.y | v.z | v.w) & 0x80808080u) == 0) return;  // sixteen ASCII bytes
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  int64_t bad = -1;
#pragma unroll 1
  for (int k = 0; k < 16 && bad < 0; ++k) {
    const int64_t i = lo + k;
    if (i >= n) break;
    const uint32_t b = (words[k >> 2] >> (8 * (k & 3))) & 0xffu;
    if (b < 0x80) continue;
    if ((b & 0xc0) == 0x80) {
      // continuation byte: some lead within the three bytes before it must reach this far
      bool covered = false;
      for (int d = 1; d <= 3 && i - d >= 0; ++d) {
        const uint32_t p = bytes[i - d];
        if ((p & 0xc0) != 0x80) {
          covered = utf8_len(p) > d;
          break;
        }
      }
      if (!covered) bad = i;
      continue;
    }
    const int len = utf8_len(b);
    if (len == 0) { bad = i; continue; }
    for (int j = 1; j < len; ++j) {
      if (i + j >= n) { bad = n; break; }
      const uint32_t c = bytes[i + j];
      uint32_t lo_ok = 0x80, hi_ok = 0xbf;
      if (j == 1) {
        if (b == 0xe0) lo_ok = 0xa0;
        else if (b == 0xed) hi_ok = 0x9f;
        else if (b == 0xf0) lo_ok = 0x90;
        else if (b == 0xf4) hi_ok = 0x8f;
      }
      if (c < lo_ok || c > hi_ok) { bad = i + j; break; }
    }
  }
  if (bad >= 0) atomicMin(first_bad, static_cast<unsigned long long>(bad));
}

// ---- field size limit (csv.field_size_limit() == 131072 code points) -----------------------------
// A field can only be too long in a row longer than the limit in bytes, so almost every thread
// returns after two loads. out[0] = first offending row, out[1] = offset of the code point that
// overflows its field.
__global__ void __launch_bounds__(256)
csv_limit_kernel(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ row_start,
                 const int64_t* __restrict__ row_end, int64_t n_rows, int64_t limit,
                 unsigned long long* out) {
  const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const int64_t a = row_start[r], e = row_end[r];
  if (e - a <= limit) return;
  uint32_t s = SR;
  int64_t flen = 0;
  for (int64_t pos = a; pos < e; ++pos) {
    const uint32_t b = bytes[pos];
    const uint32_t cls = byte_class(b);
    bool add = false;
    switch (s) {
      case SR: case ACR: case SF:
        if (cls == C_QUOTE) s = IQ;
        else if (cls == C_COMMA) { s = SF; flen = 0; }
        else { add = true; s = IF; }
        break;
      case IF:
        if (cls == C_COMMA) { s = SF; flen = 0; }
        else add = true;
        break;
      case IQ:
        if (cls == C_QUOTE) s = QQ;
        else add = true;
        break;
      default:  // QQ
        if (cls == C_QUOTE) { add = true; s = IQ; }
        else if (cls == C_COMMA) { s = SF; flen = 0; }
        else { add = true; s = IF; }
        break;
    }
    if (add && (b & 0xc0u) != 0x80u) {
      if (flen >= limit) {
        atomicMin(out, static_cast<unsigned long long>(r));
        atomicMin(out + 1, static_cast<unsigned long long>(pos));
        return;
      }
      ++flen;
    }
  }
}

// ---- rows -> fields ------------------------------------------------------------------------------
// One thread per row walks the record and either sizes it (WRITE = false: number of fields, bytes of
// field content after unquoting) or writes it (WRITE = true: content bytes back to back at
// byte_base[r], end offset of every field at field_end[field_base[r] + f]). With col >= 0 only that
// column is kept and every row yields exactly one field (empty when the row is shorter — the
// chain's restval, which map_tokenize reads as "": ops/map_tokenize.py:38).
template <bool WRITE>
__global__ void __launch_bounds__(256)
csv_rows_kernel(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ row_start,
                const int64_t* __restrict__ row_end, int64_t r0, int64_t n_rows, int32_t col,
                int64_t* __restrict__ n_fields, int64_t* __restrict__ n_bytes,
                const int64_t* __restrict__ field_base, const int64_t* __restrict__ byte_base,
                uint8_t* __restrict__ out_bytes, int64_t* __restrict__ field_end) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const int64_t a = row_start[r0 + i], e = row_end[r0 + i];
  uint32_t s = SR;
  int64_t field = 0, out = WRITE ? byte_base[i] : 0;
  const int64_t fslot = !WRITE ? 0 : col >= 0 ? i : field_base[i];
  const int64_t out0 = out;
  auto save = [&]() {
    if (col < 0) {
      if (WRITE) field_end[fslot + field] = out;
    }
    ++field;
  };
  for (int64_t pos = a; pos < e; ++pos) {
    const uint32_t b = bytes[pos];
    const uint32_t cls = byte_class(b);
    bool add = false;
    switch (s) {
      case SR: case ACR: case SF:
        if (cls == C_QUOTE) s = IQ;
        else if (cls == C_COMMA) { save(); s = SF; }
        else { add = true; s = IF; }
        break;
      case IF:
        if (cls == C_COMMA) { save(); s = SF; }
        else add = true;
        break;
      case IQ:
        if (cls == C_QUOTE) s = QQ;
        else add = true;
        break;
      default:  // QQ
        if (cls == C_QUOTE) { add = true; s = IQ; }
        else if (cls == C_COMMA) { save(); s = SF; }
        else { add = true; s = IF; }
        break;
    }
    if (add && (col < 0 || field == col)) {
      if (WRITE) out_bytes[out] = static_cast<uint8_t>(b);
      ++out;
    }
  }
  save();  // the line end (or the end of the data) closes the last field
  if (col >= 0) {
    if (WRITE) field_end[fslot] = out;
    field = 1;
  }
  if (!WRITE) {
    n_fields[i] = field;
    n_bytes[i] = out - out0;
  }
}

// ---- exclusive prefix sum of int64 (n + 1 outputs, out[n] = total) -------------------------------
constexpr int SUM_ITEMS = 8;
__global__ void __launch_bounds__(256)
sum_blocks_kernel(const int64_t* __restrict__ in, int64_t n, int64_t* __restrict__ block_sum) {
  __shared__ int64_t sm[8];
  const int64_t lo = (int64_t(blockIdx.x) * 256 + threadIdx.x) * SUM_ITEMS;
  int64_t s = 0;
  for (int k = 0; k < SUM_ITEMS; ++k)
    if (lo + k < n) s += in[lo + k];
  int64_t all = 0;
  block_sum_exclusive(s, sm, &all);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = all;
}
__global__ void __launch_bounds__(SCAN_THREADS)
scan_blocks_kernel(int64_t* __restrict__ block_sum, int64_t n_blocks, int64_t* __restrict__ total) {
  __shared__ int64_t sm[SCAN_THREADS / 32];
  const int64_t per = (n_blocks + SCAN_THREADS - 1) / SCAN_THREADS;
  const int64_t lo = min(n_blocks, int64_t(threadIdx.x) * per), hi = min(n_blocks, lo + per);
  int64_t s = 0;
  for (int64_t b = lo; b < hi; ++b) s += block_sum[b];
  int64_t all = 0;
  int64_t run = block_sum_exclusive(s, sm, &all);
  for (int64_t b = lo; b < hi; ++b) {
    const int64_t v = block_sum[b];
    block_sum[b] = run;
    run += v;
  }
  if (threadIdx.x == 0) *total = all;
}
__global__ void __launch_bounds__(256)
scan_final_kernel(const int64_t* __restrict__ in, int64_t n, const int64_t* __restrict__ block_base,
                  const int64_t* __restrict__ total, int64_t* __restrict__ out) {
  __shared__ int64_t sm[8];
  const int64_t lo = (int64_t(blockIdx.x) * 256 + threadIdx.x) * SUM_ITEMS;
  int64_t v[SUM_ITEMS];
  int64_t s = 0;
  for (int k = 0; k < SUM_ITEMS; ++k) {
    v[k] = lo + k < n ? in[lo + k] : 0;
    s += v[k];
  }
  int64_t run = block_base[blockIdx.x] + block_sum_exclusive(s, sm, nullptr);
  for (int k = 0; k < SUM_ITEMS; ++k) {
    if (lo + k < n) out[lo + k] = run;
    run += v[k];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = *total;
}

}  // namespace

// ================================================================================================
// launchers
// ================================================================================================

namespace {
struct IndexScratch {
  uint32_t* chunk_vec;
  uint8_t *chunk_state, *chunk_starts;
  uint32_t* block_vec;
  uint8_t* block_state;
  uint32_t* block_starts;
  int64_t* block_base;
  int64_t n_blocks;
};
IndexScratch carve_index(void* scratch, int64_t n_bytes) {
  IndexScratch s;
  s.n_blocks = (n_bytes + CSV_CHUNK * CSV_BLOCK - 1) / (CSV_CHUNK * CSV_BLOCK);
  if (s.n_blocks == 0) s.n_blocks = 1;
  const int64_t n_chunks = s.n_blocks * CSV_BLOCK;
  uint8_t* p = static_cast<uint8_t*>(scratch);
  auto take = [&](size_t bytes) { uint8_t* at = p; p += (bytes + 255) & ~size_t(255); return at; };
  s.block_base = reinterpret_cast<int64_t*>(take(size_t(s.n_blocks) * 8));
  s.chunk_vec = reinterpret_cast<uint32_t*>(take(size_t(n_chunks) * 4));
  s.block_vec = reinterpret_cast<uint32_t*>(take(size_t(s.n_blocks) * 4));
  s.block_starts = reinterpret_cast<uint32_t*>(take(size_t(s.n_blocks) * 4));
  s.chunk_state = take(size_t(n_chunks));
  s.chunk_starts = take(size_t(n_chunks));
  s.block_state = take(size_t(s.n_blocks));
  return s;
}
}  // namespace

size_t csv_index_scratch_bytes(int64_t n_bytes) {
  IndexScratch s = carve_index(nullptr, n_bytes);
  return reinterpret_cast<size_t>(s.block_state) + ((size_t(s.n_blocks) + 255) & ~size_t(255)) + 256;
}

int csv_count_rows(cudaStream_t st, const uint8_t* bytes, int64_t first, int64_t n, int init_state,
                   void* scratch, int64_t* d_total_rows, int32_t* d_final_state) {
  IndexScratch s = carve_index(scratch, n);
  const unsigned grid = static_cast<unsigned>(s.n_blocks);
  csv_vectors_kernel<<<grid, CSV_BLOCK, 0, st>>>(bytes, first, n, s.chunk_vec, s.block_vec);
  csv_scan_vectors_kernel<<<1, SCAN_THREADS, 0, st>>>(s.block_vec, s.n_blocks, uint32_t(init_state),
                                                      s.block_state, d_final_state);
  csv_count_kernel<<<grid, CSV_BLOCK, 0, st>>>(bytes, first, n, s.chunk_vec, s.block_state, s.chunk_state,
                                               s.chunk_starts, s.block_starts);
  csv_scan_counts_kernel<<<1, SCAN_THREADS, 0, st>>>(s.block_starts, s.n_blocks, s.block_base,
                                                     d_total_rows);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int csv_emit_rows(cudaStream_t st, const uint8_t* bytes, int64_t first, int64_t n, void* scratch,
                  int64_t* row_start, int64_t* row_end) {
  IndexScratch s = carve_index(scratch, n);
  csv_emit_kernel<<<static_cast<unsigned>(s.n_blocks), CSV_BLOCK, 0, st>>>(
      bytes, first, n, s.chunk_state, s.chunk_starts, s.block_base, row_start, row_end);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int utf8_first_invalid(cudaStream_t st, const uint8_t* bytes, int64_t n, unsigned long long* d_first_bad) {
  B2_CUDA_OK(cudaMemsetAsync(d_first_bad, 0xff, 8, st));
  if (n > 0) {
    const int64_t threads = (n + 15) / 16;
    utf8_validate_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, st>>>(bytes, n, d_first_bad);
  }
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int csv_field_limit(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                    int64_t n_rows, int64_t limit, unsigned long long* d_out2) {
  B2_CUDA_OK(cudaMemsetAsync(d_out2, 0xff, 16, st));
  if (n_rows > 0)
    csv_limit_kernel<<<static_cast<unsigned>((n_rows + 255) / 256), 256, 0, st>>>(bytes, row_start, row_end,
                                                                                n_rows, limit, d_out2);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int csv_rows_size(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                  int64_t r0, int64_t n_rows, int32_t col, int64_t* n_fields, int64_t* n_bytes) {
  if (n_rows > 0)
    csv_rows_kernel<false><<<static_cast<unsigned>((n_rows + 255) / 256), 256, 0, st>>>(
        bytes, row_start, row_end, r0, n_rows, col, n_fields, n_bytes, nullptr, nullptr, nullptr, nullptr);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int csv_rows_write(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                   int64_t r0, int64_t n_rows, int32_t col, const int64_t* field_base,
                   const int64_t* byte_base, uint8_t* out_bytes, int64_t* field_end) {
  if (n_rows > 0)
    csv_rows_kernel<true><<<static_cast<unsigned>((n_rows + 255) / 256), 256, 0, st>>>(
        bytes, row_start, row_end, r0, n_rows, col, nullptr, nullptr, field_base, byte_base, out_bytes,
        field_end);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

size_t scan_i64_scratch_bytes(int64_t n) {
  const int64_t blocks = (n + 256 * SUM_ITEMS - 1) / (256 * SUM_ITEMS) + 1;
  return size_t(blocks) * 8 + 64;
}

int exclusive_scan_i64(cudaStream_t st, const int64_t* in, int64_t n, int64_t* out, void* scratch) {
  int64_t* total = static_cast<int64_t*>(scratch);
  int64_t* block_sum = total + 8;
  const int64_t blocks = (n + 256 * SUM_ITEMS - 1) / (256 * SUM_ITEMS);
  if (n <= 0) {
    B2_CUDA_OK(cudaMemsetAsync(out, 0, 8, st));
    return B2_OK;
  }
  sum_blocks_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(in, n, block_sum);
  scan_blocks_kernel<<<1, SCAN_THREADS, 0, st>>>(block_sum, blocks, total);
  scan_final_kernel<<<static_cast<unsigned>(blocks), 256, 0, st>>>(in, n, block_sum, total, out);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace b2
