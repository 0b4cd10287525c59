// b2_csv.cu — CSV record splitter and field extractor for the `read_csv_shard` op and the
// csv -> tokenize -> classify chain (reference: ops/csv_shard.py:9-26, which drives Python's
// csv.DictReader over a file opened with newline="" and encoding="utf-8").
//
// What has to be reproduced is the state machine of CPython's csv reader with its default dialect
// (delimiter ',', quotechar '"', doublequote on, no escapechar, strict off), fed by a line iterator
// that ends a physical line after "\n", "\r" or "\r\n". On raw bytes that is a five-state automaton
// over four byte classes (quote, comma, line end, anything else):
//
//   SR  start of record             SF  start of field (after ',')
//   IF  inside an unquoted field    IQ  inside a quoted field
//   QQ  just saw a quote inside a quoted field
//
// A record that starts (first byte that is not a line end, seen in SR) is a row; line ends seen in
// SR are the reader's empty records, which DictReader drops — which is also why "\r\n" needs no
// state of its own here: the '\n' is just one more line end seen in SR. Quotes only open a field at
// its start, a quote in the middle of an unquoted field is data, and after a closing quote any
// other byte continues the field unquoted — none of which a quote-parity scan gets right, hence the
// automaton.
//
// The file is cut into 64-byte chunks, one per thread. Pass 1 (`csv_classify_kernel`) reads the
// bytes once and reduces every chunk to three 64-bit masks (quotes, commas, line ends); every later
// step works on the masks and visits structural bytes only, whatever lies between two of them is
// one application of the "other byte" transition. SR, SF and QQ have the same successors, so a
// chunk's effect on the state is a function of three classes of start state ({SR,SF,QQ}, IF, IQ):
// pass 1 runs the three in parallel, packs the three end states into 9 bits and composes them per
// CTA; a single-CTA scan turns the CTA functions into the true state at every CTA boundary. Pass 2
// (`csv_count_kernel`) re-walks the masks from the now known states and counts row starts; after a
// scan of the counts, pass 3 (`csv_emit_kernel`) writes the byte offset of every row start and row
// end. Pass 1 also validates the bytes as UTF-8 while it has them in registers. HBM traffic per
// chunk: 64 B read + 26 B written in pass 1, 26 + 2 and 28 B read in passes 2 and 3.
#include "b2_kernels.h"

namespace b2 {
namespace {

constexpr int CSV_CHUNK = 64;   // bytes per thread
constexpr int CSV_BLOCK = 256;  // threads per block: 16 KiB of file per block
constexpr int SCAN_THREADS = 1024;

enum : uint32_t { SR = 0, SF = 2, IF = 3, IQ = 4, QQ = 5 };  // 1 is unused (b2_csv_info.end_state > 1
                                                             // keeps meaning "inside a record")
enum : uint32_t { C_OTHER = 0, C_QUOTE = 1, C_COMMA = 2, C_NL = 3 };

constexpr uint32_t pack6(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f) {
  return a | (b << 3) | (c << 6) | (d << 9) | (e << 12) | (f << 15);
}
// next state, indexed by the current state (3 bits each), one word per byte class
constexpr uint32_t F_OTHER = pack6(IF, IF, IF, IF, IQ, IF);
constexpr uint32_t F_QUOTE = pack6(IQ, IQ, IQ, IF, QQ, IQ);
constexpr uint32_t F_COMMA = pack6(SF, SF, SF, SF, IQ, SF);
constexpr uint32_t F_NL = pack6(SR, SR, SR, SR, IQ, SR);

__device__ __forceinline__ uint32_t byte_class(uint32_t b) {
  return b == '"' ? C_QUOTE : b == ',' ? C_COMMA : (b == '\n' || b == '\r') ? C_NL : C_OTHER;
}
__device__ __forceinline__ uint32_t apply(uint32_t fn, uint32_t state) { return (fn >> (3 * state)) & 7u; }

// A chunk's (or any byte range's) effect on the state, "fn3": bits 0-2 end state for a start in
// {SR, SF, QQ}, bits 3-5 for a start in IF, bits 6-8 for a start in IQ; FN3_IDENT marks the empty
// range, the one function the three classes cannot express.
constexpr uint32_t FN3_IDENT = 1u << 9;
constexpr uint32_t STATE_CLASS = (1u << (2 * IF)) | (2u << (2 * IQ));  // state -> 0, 1 (IF), 2 (IQ)
__device__ __forceinline__ uint32_t apply3(uint32_t fn, uint32_t state) {
  const uint32_t cls = (STATE_CLASS >> (2 * state)) & 3u;
  return (fn & FN3_IDENT) ? state : (fn >> (3 * cls)) & 7u;
}
// (a then b)
__device__ __forceinline__ uint32_t compose3(uint32_t a, uint32_t b) {
  const uint32_t r = apply3(b, a & 7u) | (apply3(b, (a >> 3) & 7u) << 3) | (apply3(b, (a >> 6) & 7u) << 6);
  return (a & FN3_IDENT) ? b : (b & FN3_IDENT) ? a : r;
}

// 4-bit mask of the bytes of w equal to the byte replicated in pat (exact, no carries across bytes)
__device__ __forceinline__ uint32_t eq_mask4(uint32_t w, uint32_t pat) {
  const uint32_t x = w ^ pat;
  const uint32_t t = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);  // 0x80 where x == 0
  return ((t >> 7) * 0x00204081u) >> 21 & 0xfu;
}

struct ChunkMasks {
  unsigned long long quote, comma, nl;
};

// bits [lo_bit, hi_bit) of a 64-bit mask
__device__ __forceinline__ unsigned long long range_mask(int lo_bit, int hi_bit) {
  const unsigned long long hi = hi_bit >= 64 ? ~0ull : ((1ull << hi_bit) - 1ull);
  const unsigned long long lo = lo_bit >= 64 ? ~0ull : ((1ull << lo_bit) - 1ull);
  return hi & ~lo;
}

// Walk one chunk's structural bytes from a known state. vlo / vhi: first and one-past-last valid
// byte of the chunk (chunk-relative, 0..64; bytes outside belong to the header or lie beyond the
// data). on_start(p) / on_end(p): a row starts at / is closed by the line end at chunk byte p.
template <typename OnStart, typename OnEnd>
__device__ __forceinline__ uint32_t walk_masks(const ChunkMasks& mk, int vlo, int vhi, uint32_t state,
                                               OnStart&& on_start, OnEnd&& on_end) {
  if (vhi <= vlo) return state;
  unsigned long long m = (mk.quote | mk.comma | mk.nl) & range_mask(vlo, vhi);
  int prev = vlo - 1;
  while (m) {
    const int p = __ffsll(static_cast<long long>(m)) - 1;
    m &= m - 1;
    if (p > prev + 1) {  // other bytes in between
      if (state == SR) on_start(prev + 1);
      state = apply(F_OTHER, state);
    }
    const bool is_nl = (mk.nl >> p) & 1ull;
    const uint32_t fn = is_nl ? F_NL : ((mk.quote >> p) & 1ull) ? F_QUOTE : F_COMMA;
    if (state == SR) {
      if (!is_nl) on_start(p);
    } else if (state != IQ && is_nl) {
      on_end(p);
    }
    state = apply(fn, state);
    prev = p;
  }
  if (vhi - 1 > prev) {
    if (state == SR) on_start(prev + 1);
    state = apply(F_OTHER, state);
  }
  return state;
}

// Ordered composition over the block: thread 0 returns the composition of all threads' v.
__device__ uint32_t block_compose_all(uint32_t v, uint32_t* smem /* CSV_BLOCK / 32 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t right = __shfl_down_sync(0xffffffffu, v, off);
    if ((lane & (2 * off - 1)) == 0) v = compose3(v, right);
  }
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  uint32_t total = FN3_IDENT;
  if (threadIdx.x == 0)
    for (int w = 0; w < CSV_BLOCK / 32; ++w) total = compose3(total, smem[w]);
  return total;
}
// Ordered exclusive scan over the block: composition of the functions of all threads before this one.
__device__ uint32_t block_compose_exclusive(uint32_t v, uint32_t* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t left = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl = compose3(left, incl);
  }
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  uint32_t before = FN3_IDENT;
  for (int w = 0; w < warp; ++w) before = compose3(before, smem[w]);
  uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = FN3_IDENT;
  return compose3(before, excl);
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

// ---- UTF-8 validation (the reference decodes the file strictly: ops/csv_shard.py:17) -------------
// The offset reported is the smallest one at which a strict decoder holds the evidence of an
// error: the stray or illegal byte itself, the continuation byte that breaks a sequence, or n
// when the data ends inside a sequence.
__device__ __forceinline__ int utf8_len(uint32_t b) {
  if (b < 0x80) return 1;
  if (b < 0xc2) return 0;  // continuation byte or overlong 2-byte lead
  if (b < 0xe0) return 2;
  if (b < 0xf0) return 3;
  if (b < 0xf5) return 4;
  return 0;
}
// i: offset of a byte >= 0x80. Returns the offset of the evidence, or -1 if byte i is fine.
__device__ int64_t utf8_check_byte(const uint8_t* __restrict__ bytes, int64_t i, int64_t n) {
  const uint32_t b = bytes[i];
  if ((b & 0xc0) == 0x80) {
    // continuation byte: some lead within the three bytes before it must reach this far
    for (int d = 1; d <= 3 && i - d >= 0; ++d) {
      const uint32_t p = bytes[i - d];
      if ((p & 0xc0) != 0x80) return utf8_len(p) > d ? -1 : i;
    }
    return i;
  }
  const int len = utf8_len(b);
  if (len == 0) return i;
  for (int j = 1; j < len; ++j) {
    if (i + j >= n) return n;
    const uint32_t c = bytes[i + j];
    uint32_t lo_ok = 0x80, hi_ok = 0xbf;
    if (j == 1) {
      if (b == 0xe0) lo_ok = 0xa0;
      else if (b == 0xed) hi_ok = 0x9f;
      else if (b == 0xf0) lo_ok = 0x90;
      else if (b == 0xf4) hi_ok = 0x8f;
    }
    if (c < lo_ok || c > hi_ok) return i + j;
  }
  return -1;
}

// ---- pass 1: bytes -> masks, chunk functions composed per block, UTF-8 check ----------------------
// Validation covers every byte of the buffer (the header too); the masks only bytes [first, n).
__global__ void __launch_bounds__(CSV_BLOCK)
csv_classify_kernel(const uint8_t* __restrict__ bytes, int64_t first, int64_t n,
                    unsigned long long* __restrict__ mask_quote, unsigned long long* __restrict__ mask_comma,
                    unsigned long long* __restrict__ mask_nl, uint16_t* __restrict__ chunk_fn,
                    uint32_t* __restrict__ block_fn, unsigned long long* __restrict__ first_bad) {
  __shared__ uint32_t sm[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  ChunkMasks mk{0ull, 0ull, 0ull};
  unsigned long long high = 0ull;  // bytes >= 0x80
  if (lo < n) {
    const uint4* src = reinterpret_cast<const uint4*>(bytes + lo);
#pragma unroll
    for (int q = 0; q < CSV_CHUNK / 16; ++q) {
      if (lo + q * 16 < n) {
        const uint4 ld = __ldg(src + q);
        const uint32_t words[4] = {ld.x, ld.y, ld.z, ld.w};
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const int sh = q * 16 + w * 4;
          const uint32_t word = words[w];
          mk.quote |= static_cast<unsigned long long>(eq_mask4(word, 0x22222222u)) << sh;
          mk.comma |= static_cast<unsigned long long>(eq_mask4(word, 0x2c2c2c2cu)) << sh;
          mk.nl |= static_cast<unsigned long long>(eq_mask4(word, 0x0a0a0a0au) | eq_mask4(word, 0x0d0d0d0du)) << sh;
          high |= static_cast<unsigned long long>((((word & 0x80808080u) >> 7) * 0x00204081u) >> 21 & 0xfu) << sh;
        }
      }
    }
  }
  const int vlo = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), first - lo)));
  const int vhi = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), n - lo)));
  const unsigned long long valid = vhi > vlo ? range_mask(vlo, vhi) : 0ull;
  mk.quote &= valid;
  mk.comma &= valid;
  mk.nl &= valid;

  // the chunk as a function of the three start classes
  uint32_t fn = FN3_IDENT;
  if (vhi > vlo) {
    uint32_t s0 = SR, s1 = IF, s2 = IQ;
    unsigned long long m = mk.quote | mk.comma | mk.nl;
    int prev = vlo - 1;
    while (m) {
      const int p = __ffsll(static_cast<long long>(m)) - 1;
      m &= m - 1;
      if (p > prev + 1) {
        s0 = apply(F_OTHER, s0);
        s1 = apply(F_OTHER, s1);
        s2 = apply(F_OTHER, s2);
      }
      const uint32_t f = ((mk.nl >> p) & 1ull) ? F_NL : ((mk.quote >> p) & 1ull) ? F_QUOTE : F_COMMA;
      s0 = apply(f, s0);
      s1 = apply(f, s1);
      s2 = apply(f, s2);
      prev = p;
    }
    if (vhi - 1 > prev) {
      s0 = apply(F_OTHER, s0);
      s1 = apply(F_OTHER, s1);
      s2 = apply(F_OTHER, s2);
    }
    fn = s0 | (s1 << 3) | (s2 << 6);
  }
  mask_quote[chunk] = mk.quote;
  mask_comma[chunk] = mk.comma;
  mask_nl[chunk] = mk.nl;
  chunk_fn[chunk] = static_cast<uint16_t>(fn);

  // UTF-8: only bytes with the top bit set need a look
  if (lo < n) high &= range_mask(0, static_cast<int>(min(int64_t(CSV_CHUNK), n - lo)));
  int64_t bad = -1;
  while (high && bad < 0) {
    const int p = __ffsll(static_cast<long long>(high)) - 1;
    high &= high - 1;
    bad = utf8_check_byte(bytes, lo + p, n);
  }
  if (bad >= 0) atomicMin(first_bad, static_cast<unsigned long long>(bad));

  const uint32_t total = block_compose_all(fn, sm);
  if (threadIdx.x == 0) block_fn[blockIdx.x] = total;
}

// ---- single-CTA scans over the per-block results ------------------------------------------------
// Warp w owns the contiguous range [w*per, (w+1)*per) of blocks and walks it 32 entries at a time
// (coalesced loads, shuffle scan, running carry); the 32 warp totals are then combined and each
// warp walks its range once more to write the results.
__global__ void __launch_bounds__(SCAN_THREADS)
csv_scan_fn_kernel(const uint32_t* __restrict__ block_fn, int64_t n_blocks, uint32_t init_state,
                   uint8_t* __restrict__ block_state, int32_t* __restrict__ final_state) {
  __shared__ uint32_t warp_total[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t per = ((n_blocks + 31) / 32 + 31) / 32 * 32;  // per warp, a multiple of 32
  const int64_t lo = min(n_blocks, int64_t(warp) * per), hi = min(n_blocks, lo + per);
  uint32_t carry = FN3_IDENT;
  for (int64_t base = lo; base < hi; base += 32) {
    uint32_t v = base + lane < hi ? block_fn[base + lane] : FN3_IDENT;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t left = __shfl_up_sync(0xffffffffu, v, off);
      if (lane >= off) v = compose3(left, v);
    }
    carry = compose3(carry, __shfl_sync(0xffffffffu, v, 31));
  }
  if (lane == 0) warp_total[warp] = carry;
  __syncthreads();
  uint32_t state = init_state;
  for (int w = 0; w < warp; ++w) state = apply3(warp_total[w], state);
  // second walk: state before every block
  for (int64_t base = lo; base < hi; base += 32) {
    const uint32_t mine = base + lane < hi ? block_fn[base + lane] : FN3_IDENT;
    uint32_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t left = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl = compose3(left, incl);
    }
    uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = FN3_IDENT;
    if (base + lane < hi) block_state[base + lane] = static_cast<uint8_t>(apply3(excl, state));
    state = apply3(__shfl_sync(0xffffffffu, incl, 31), state);
  }
  if (threadIdx.x == SCAN_THREADS - 1) *final_state = static_cast<int32_t>(state);
}

__global__ void __launch_bounds__(SCAN_THREADS)
csv_scan_counts_kernel(const uint32_t* __restrict__ block_count, int64_t n_blocks,
                       int64_t* __restrict__ block_base, int64_t* __restrict__ total) {
  __shared__ int64_t warp_total[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t per = ((n_blocks + 31) / 32 + 31) / 32 * 32;
  const int64_t lo = min(n_blocks, int64_t(warp) * per), hi = min(n_blocks, lo + per);
  int64_t sum = 0;
  for (int64_t base = lo + lane; base < hi; base += 32) sum += block_count[base];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  if (lane == 0) warp_total[warp] = sum;
  __syncthreads();
  int64_t run = 0, all = 0;
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    if (w < warp) run += warp_total[w];
    all += warp_total[w];
  }
  for (int64_t base = lo; base < hi; base += 32) {
    const int64_t mine = base + lane < hi ? block_count[base + lane] : 0;
    int64_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int64_t left = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += left;
    }
    if (base + lane < hi) block_base[base + lane] = run + incl - mine;
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (threadIdx.x == 0) *total = all;
}

// ---- pass 2: true start state of every chunk, row starts per chunk ------------------------------
__global__ void __launch_bounds__(CSV_BLOCK)
csv_count_kernel(int64_t first, int64_t n, const unsigned long long* __restrict__ mask_quote,
                 const unsigned long long* __restrict__ mask_comma,
                 const unsigned long long* __restrict__ mask_nl, const uint16_t* __restrict__ chunk_fn,
                 const uint8_t* __restrict__ block_state, uint8_t* __restrict__ chunk_state,
                 uint8_t* __restrict__ chunk_starts, uint32_t* __restrict__ block_starts) {
  __shared__ uint32_t sm[CSV_BLOCK / 32];
  __shared__ int64_t sm64[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  const uint32_t prefix = block_compose_exclusive(chunk_fn[chunk], sm);
  const uint32_t state = apply3(prefix, block_state[blockIdx.x]);
  const ChunkMasks mk{mask_quote[chunk], mask_comma[chunk], mask_nl[chunk]};
  const int vlo = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), first - lo)));
  const int vhi = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), n - lo)));
  int starts = 0;
  walk_masks(mk, vlo, vhi, state, [&](int) { ++starts; }, [](int) {});
  chunk_state[chunk] = static_cast<uint8_t>(state);
  chunk_starts[chunk] = static_cast<uint8_t>(starts);
  int64_t all = 0;
  block_sum_exclusive(starts, sm64, &all);
  if (threadIdx.x == 0) block_starts[blockIdx.x] = static_cast<uint32_t>(all);
}

// ---- pass 3: byte offsets of row starts and row ends -------------------------------------------
__global__ void __launch_bounds__(CSV_BLOCK)
csv_emit_kernel(int64_t first, int64_t n, const unsigned long long* __restrict__ mask_quote,
                const unsigned long long* __restrict__ mask_comma,
                const unsigned long long* __restrict__ mask_nl, const uint8_t* __restrict__ chunk_state,
                const uint8_t* __restrict__ chunk_starts, const int64_t* __restrict__ block_base,
                int64_t* __restrict__ row_start, int64_t* __restrict__ row_end) {
  __shared__ int64_t sm64[CSV_BLOCK / 32];
  const int64_t chunk = int64_t(blockIdx.x) * CSV_BLOCK + threadIdx.x;
  const int64_t lo = chunk * CSV_CHUNK;
  int64_t row = block_base[blockIdx.x] + block_sum_exclusive(chunk_starts[chunk], sm64, nullptr);
  const ChunkMasks mk{mask_quote[chunk], mask_comma[chunk], mask_nl[chunk]};
  const int vlo = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), first - lo)));
  const int vhi = static_cast<int>(max(int64_t(0), min(int64_t(CSV_CHUNK), n - lo)));
  walk_masks(
      mk, vlo, vhi, chunk_state[chunk], [&](int p) { row_start[row++] = lo + p; },
      [&](int p) { row_end[row - 1] = lo + p; });
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
      case SR: case SF:
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
      case SR: case SF:
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
  unsigned long long *mask_quote, *mask_comma, *mask_nl;
  int64_t* block_base;
  uint32_t *block_fn, *block_starts;
  uint16_t* chunk_fn;
  uint8_t *chunk_state, *chunk_starts, *block_state;
  int64_t n_blocks;
  size_t bytes;
};
IndexScratch carve_index(void* scratch, int64_t n_bytes) {
  IndexScratch s;
  s.n_blocks = (n_bytes + CSV_CHUNK * CSV_BLOCK - 1) / (CSV_CHUNK * CSV_BLOCK);
  if (s.n_blocks == 0) s.n_blocks = 1;
  const int64_t n_chunks = s.n_blocks * CSV_BLOCK;
  uint8_t* const base = static_cast<uint8_t*>(scratch);
  uint8_t* p = base;
  auto take = [&](size_t bytes) { uint8_t* at = p; p += (bytes + 255) & ~size_t(255); return at; };
  s.mask_quote = reinterpret_cast<unsigned long long*>(take(size_t(n_chunks) * 8));
  s.mask_comma = reinterpret_cast<unsigned long long*>(take(size_t(n_chunks) * 8));
  s.mask_nl = reinterpret_cast<unsigned long long*>(take(size_t(n_chunks) * 8));
  s.block_base = reinterpret_cast<int64_t*>(take(size_t(s.n_blocks) * 8));
  s.block_fn = reinterpret_cast<uint32_t*>(take(size_t(s.n_blocks) * 4));
  s.block_starts = reinterpret_cast<uint32_t*>(take(size_t(s.n_blocks) * 4));
  s.chunk_fn = reinterpret_cast<uint16_t*>(take(size_t(n_chunks) * 2));
  s.chunk_state = take(size_t(n_chunks));
  s.chunk_starts = take(size_t(n_chunks));
  s.block_state = take(size_t(s.n_blocks));
  s.bytes = static_cast<size_t>(p - base);
  return s;
}
}  // namespace

size_t csv_index_scratch_bytes(int64_t n_bytes) { return carve_index(nullptr, n_bytes).bytes + 256; }

int csv_count_rows(cudaStream_t st, const uint8_t* bytes, int64_t first, int64_t n, void* scratch,
                   int64_t* d_total_rows, int32_t* d_final_state, unsigned long long* d_first_bad) {
  IndexScratch s = carve_index(scratch, n);
  const unsigned grid = static_cast<unsigned>(s.n_blocks);
  B2_CUDA_OK(cudaMemsetAsync(d_first_bad, 0xff, 8, st));
  csv_classify_kernel<<<grid, CSV_BLOCK, 0, st>>>(bytes, first, n, s.mask_quote, s.mask_comma, s.mask_nl,
                                                  s.chunk_fn, s.block_fn, d_first_bad);
  csv_scan_fn_kernel<<<1, SCAN_THREADS, 0, st>>>(s.block_fn, s.n_blocks, SR, s.block_state, d_final_state);
  csv_count_kernel<<<grid, CSV_BLOCK, 0, st>>>(first, n, s.mask_quote, s.mask_comma, s.mask_nl, s.chunk_fn,
                                               s.block_state, s.chunk_state, s.chunk_starts, s.block_starts);
  csv_scan_counts_kernel<<<1, SCAN_THREADS, 0, st>>>(s.block_starts, s.n_blocks, s.block_base,
                                                     d_total_rows);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int csv_emit_rows(cudaStream_t st, int64_t first, int64_t n, void* scratch, int64_t* row_start,
                  int64_t* row_end) {
  IndexScratch s = carve_index(scratch, n);
  csv_emit_kernel<<<static_cast<unsigned>(s.n_blocks), CSV_BLOCK, 0, st>>>(
      first, n, s.mask_quote, s.mask_comma, s.mask_nl, s.chunk_state, s.chunk_starts, s.block_base,
      row_start, row_end);
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
