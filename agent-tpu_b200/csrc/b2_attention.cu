// b2_attention.cu — masked self-attention for the encoder, head dim 64, on tcgen05 + TMA.
//
// Persistent kernels, two CTAs per SM. A CTA walks its work items through a two-stage ring of
// {Q, K, V} tiles in shared memory and 256 columns of tensor memory:
//   S = Q·Kᵀ            tcgen05.mma 128 x n x 64  (Q, K: K-major 128-byte-swizzled TMA tiles)
//   P = exp2((S-m)·c)    two threads per query row read S from TMEM; keys >= len are masked out
//   O = P·V              tcgen05.mma 128 x 64 x n  (V: MN-major TMA tile, no transpose)
// with n = the block's valid keys rounded up to 32: the TMA boxes (four tensor maps with 32, 64, 96
// and 128 rows), the UMMA N of Q·Kᵀ and the K extent of P·V all shrink with the sequence. Rows of
// the tiles beyond the box hold stale shared memory; no MMA whose result is kept reads them.
//
// S = 128 kernel (attention_persist_kernel), per item i:
//   * P(i) never touches shared memory: the softmax threads write it as packed bf16 over the S
//     columns they have just consumed and P·V takes its A operand from tensor memory. That halves
//     the shared-memory traffic of P·V (only V is streamed), removes the row-strided P stores and
//     the proxy fence from the softmax chain, and frees the Q and K tiles the moment Q·Kᵀ(i) is
//     done, so {Q, K}(i+2) is requested a whole softmax earlier than V(i+2);
//   * S has one buffer and O two: Q·Kᵀ(i+1) is issued right behind P·V(i) and the softmax threads
//     normalise and store O(i-1) while the tensor core works on both, instead of idling until P·V(i)
//     has finished and then again until Q·Kᵀ(i+1) has (dev_check timeline: 1.0 + 0.7 us of the
//     former 3.6 us per item were those two waits).
//     Variants measured and dropped (full-length batch, 0.153 ms as built): P in columns of its own
//     with Q·Kᵀ(i+1) issued ahead of P·V(i) and a single O buffer — 0.158 ms, because P·V(i) then
//     waits for the O(i-1) read-out and the next P write waits for P·V(i); one CTA per SM — 0.207 ms.
//     With 256 of the SM's 512 TMEM columns per CTA there is room for two item buffers, and a
//     bubble-free schedule needs three; the short MMAs (32-64 cycles of math each) also run at about
//     half their floor rate when issued in bursts of 12 per item.
//   MMA thread / TMA thread: two single-thread roles in two warps, because one thread doing both
//     spent 0.8 us per item issuing (dev_check timeline); per-item (length, first row) are looked up
//     once into shared memory: a global load in such a dependent chain costs ~0.5 us each time.
//     Tiles two items further ahead are prefetched into L2.
//   softmax warps: eight per CTA, two threads per query row (keys 0-63 / 64-127, then output columns
//     0-31 / 32-63); the row maximum and denominator are exchanged through shared memory. This part
//     is a chain of TMEM loads and MUFU.EX2: it needs warps, not width, to hide its latency.
//
// attention_persist_kernel  S = 128: one item = (sequence, head), one key block.
// attention_stream_kernel   S = 256..512: one ring step = (sequence, head, query block, key block);
//                           Q is re-fetched with every key block (an L2 hit), P goes through shared
//                           memory over the dead Q|K tiles, and the softmax threads carry
//                           the running maximum, denominator and f32 output row across the key blocks
//                           (online softmax); blocks wholly beyond the sequence are skipped.
//
// No reference counterpart: the reference's model is an opaque TFLite graph
// (ops/map_classify_tpu.py:71-72). Semantics follow BertSelfAttention (scores/sqrt(d) + additive
// key mask, softmax in f32, context = probs·V).
#include <cstdlib>

#include "b2_kernels.h"

namespace b2 {

namespace {

constexpr int ATT_THREADS = 320;  // warps 0-7: softmax (2 threads per query row); warp 8: MMA issue; warp 9: TMA
constexpr int ATT_SOFTMAX_THREADS = 256;
constexpr int QB = 128;  // queries per item / step
constexpr int KB = 128;  // keys per block
constexpr int HD = 64;   // head dim
constexpr uint32_t TILE_BYTES = 128 * 64 * 2;  // 16 KiB: a [128 x 64] bf16 tile
constexpr uint32_t STAGE_BYTES = 3 * TILE_BYTES;
constexpr int PERSIST_MAX_ITEMS = 256;  // items one CTA of the S=128 kernel may walk (grid is sized for it)
constexpr uint32_t OFF_BARS = 2 * STAGE_BYTES;
constexpr uint32_t OFF_RED = OFF_BARS + 128;          // float[4][128]: max[2][128], sum[2][128]
constexpr uint32_t OFF_META = OFF_RED + 4 * 128 * 4;  // int2[PERSIST_MAX_ITEMS]
constexpr uint32_t ATT_SMEM = 1024 + OFF_META + PERSIST_MAX_ITEMS * 8;

unsigned long long* g_att_trace = nullptr;

struct AttMaps {
  CUtensorMap m[4];  // box rows 32, 64, 96, 128 over the same [rows, 3H] buffer
};

enum { QKV_FULL = 0, S_FULL = 1, P_FULL = 2, O_FULL = 3, FREE = 4 };
// the S = 128 kernel loads {Q, K} and V separately and has no FREE barrier
enum { QK_FULL = QKV_FULL, V_FULL = FREE };

__device__ __forceinline__ void softmax_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Row maximum over this thread's half row: up to 64 keys = two 32-column TMEM chunks, one register
// buffer (288 threads x 2 CTAs/SM leave 96 registers per thread; two buffers in flight spilled).
// A chunk wholly inside the sequence uses FMNMX3 on key pairs.
__device__ __forceinline__ float softmax_max64(uint32_t tS, int kv) {
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (c * 32 < kv) {
      uint32_t v[32];
      tmem_ld_32x32(tS + c * 32, v);
      tmem_ld_wait();
      if (c * 32 + 32 <= kv) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) mx = fmaxf(mx, fmaxf(__uint_as_float(v[j]), __uint_as_float(v[j + 1])));
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c * 32 + j < kv) mx = fmaxf(mx, __uint_as_float(v[j]));
      }
    }
  }
  return mx;
}

// exp2((s - m)·c) for one 32-key chunk -> 16 registers of packed bf16 P; returns the chunk's sum.
// Unmasked chunks: packed FFMA2 / FADD2, two MUFU.EX2 and one bf16x2 convert per key pair.
__device__ __forceinline__ float softmax_exp32_regs(const uint32_t (&v)[32], int kv_chunk, float sl2, float moff,
                                                    uint32_t (&packed)[16]) {
  float sum;
  if (kv_chunk >= 32) {
    const uint64_t scale2 = f32x2(sl2, sl2), shift2 = f32x2(-moff, -moff);
    uint64_t sum2 = f32x2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      float p0, p1;
      f32x2_split(ffma2(f32x2_bits(v[j], v[j + 1]), scale2, shift2), p0, p1);
      p0 = ex2_approx(p0);
      p1 = ex2_approx(p1);
      sum2 = fadd2(sum2, f32x2(p0, p1));
      packed[j >> 1] = pack_bf16x2(p0, p1);
    }
    float s0, s1;
    f32x2_split(sum2, s0, s1);
    sum = s0 + s1;
  } else {
    sum = 0.f;
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      const float p0 = (j < kv_chunk) ? ex2_approx(fmaf(__uint_as_float(v[j]), sl2, -moff)) : 0.f;
      const float p1 = (j + 1 < kv_chunk) ? ex2_approx(fmaf(__uint_as_float(v[j + 1]), sl2, -moff)) : 0.f;
      sum += p0 + p1;
      packed[j >> 1] = pack_bf16x2(p0, p1);
    }
  }
  return sum;
}

// ... -> bf16 P in the swizzled shared-memory tile row (the streaming kernel's A operand)
__device__ __forceinline__ float softmax_exp32(const uint32_t (&v)[32], int kv_chunk, float sl2, float moff,
                                               uint8_t* tile_row, int c, int r) {
  uint32_t packed[16];
  const float sum = softmax_exp32_regs(v, kv_chunk, sl2, moff, packed);
  // chunk c (0/1) of this 64-key tile -> 16-byte slots (4c + j) ^ (r & 7) of the 128-byte row
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int slot = ((c * 4 + j) ^ (r & 7));
    *reinterpret_cast<uint4*>(tile_row + slot * 16) =
        make_uint4(packed[j * 4], packed[j * 4 + 1], packed[j * 4 + 2], packed[j * 4 + 3]);
  }
  return sum;
}

// P for this thread's half row, written over the S columns it has just consumed: chunk c of the half
// row (S columns [32c, 32c+32) of the half) becomes 16 columns of packed bf16 at [16c, 16c+16) of the
// half — the layout tcgen05.mma takes for an A operand in tensor memory.
__device__ __forceinline__ float softmax_exp64_tmem(uint32_t tS_half, int kv, float sl2, float m_new) {
  const float moff = m_new * sl2;
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (c * 32 < kv) {
      uint32_t v[32];
      tmem_ld_32x32(tS_half + c * 32, v);
      tmem_ld_wait();
      uint32_t packed[16];
      sum += softmax_exp32_regs(v, kv - c * 32, sl2, moff, packed);
      tmem_st_32x16(tS_half + c * 16, packed);
    }
  }
  tmem_st_wait();
  return sum;
}

// P for this thread's half row (up to 64 keys) given the row maximum; returns the half row's sum.
__device__ __forceinline__ float softmax_exp64(uint32_t tS, int kv, float sl2, float m_new, uint8_t* tile_row,
                                               int r) {
  const float moff = m_new * sl2;
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    if (c * 32 < kv) {
      uint32_t v[32];
      tmem_ld_32x32(tS + c * 32, v);
      tmem_ld_wait();
      sum += softmax_exp32(v, kv - c * 32, sl2, moff, tile_row, c, r);
    }
  }
  return sum;
}

// ================================================================================================
// S = 128: one item = (sequence, head)
// ================================================================================================
__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_persist_kernel(const __grid_constant__ AttMaps maps, __nv_bfloat16* __restrict__ out,
                         const int32_t* __restrict__ lens, int S, int hidden, int heads, int n_items,
                         const int32_t* __restrict__ cu, unsigned long long* __restrict__ trace) {
  // developer timeline (dev_check "atttrace"): CTA 0 records %globaltimer at each pipeline event of
  // its first 64 items: [item][0..1] MMA thread (P seen, P·V and next Q·Kᵀ issued), [item][3] TMA thread
  // (next-next loads issued),
  // [item][4..7] softmax thread 0 (S seen, P arrived, O seen, row stored)
  auto stamp = [&](int item, int slot) {
    if (trace && blockIdx.x == 0 && item < 64) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      trace[item * 8 + slot] = t;
    }
  };
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + OFF_BARS;
  auto bar = [&](int which, int s) { return bars + 8u * (which * 2 + s); };
  const uint32_t tmem_slot = bars + 8u * 10;
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_aligned + OFF_BARS + 80);
  float* red = reinterpret_cast<float*>(smem_aligned + OFF_RED);
  int2* meta = reinterpret_cast<int2*>(smem_aligned + OFF_META);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int my_items = (n_items - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);
  for (int i = threadIdx.x; i < my_items; i += blockDim.x) {
    const int idx = blockIdx.x + i * gridDim.x;
    const int b = idx / heads;
    int len = lens[b];
    len = len < 1 ? 1 : (len > S ? S : len);
    meta[i] = make_int2(len, cu ? cu[b] : b * S);
  }

  if (threadIdx.x == ATT_SOFTMAX_THREADS) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(QK_FULL, s), 1);
      mbar_init(bar(V_FULL, s), 1);
      mbar_init(bar(S_FULL, s), 1);
      mbar_init(bar(P_FULL, s), ATT_SOFTMAX_THREADS);
      mbar_init(bar(O_FULL, s), 1);
    }
    fence_barrier_init();
  }
  if (warp == 8) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 9) {
    // ===== TMA thread =====
    if (lane == 0) {
      // {Q, K} and V of an item travel separately: the Q and K tiles are dead as soon as Q·Kᵀ has
      // completed (P lives in tensor memory), the V tile when P·V has
      auto load_qk = [&](int i) {
        const int idx = blockIdx.x + i * gridDim.x;
        const int h = idx % heads;
        const int2 md = meta[i];
        const int box = (md.x + 31) >> 5;
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES;
        mbar_expect_tx(bar(QK_FULL, s), 2 * box * 32 * 128);
        tma_load_2d(st, &maps.m[box - 1], bar(QK_FULL, s), h * HD, md.y);
        tma_load_2d(st + TILE_BYTES, &maps.m[box - 1], bar(QK_FULL, s), hidden + h * HD, md.y);
      };
      auto load_v = [&](int i) {
        const int idx = blockIdx.x + i * gridDim.x;
        const int h = idx % heads;
        const int2 md = meta[i];
        const int box = (md.x + 31) >> 5;
        const int s = i & 1;
        mbar_expect_tx(bar(V_FULL, s), box * 32 * 128);
        tma_load_2d(base + s * STAGE_BYTES + 2 * TILE_BYTES, &maps.m[box - 1], bar(V_FULL, s),
                    2 * hidden + h * HD, md.y);
      };
      // pull item i's tiles from DRAM into L2 ahead of its TMA load (no shared memory needed)
      auto prefetch_item = [&](int i) {
        const int idx = blockIdx.x + i * gridDim.x;
        const int h = idx % heads;
        const int2 md = meta[i];
        const int box = (md.x + 31) >> 5;
        tma_prefetch_2d(&maps.m[box - 1], h * HD, md.y);
        tma_prefetch_2d(&maps.m[box - 1], hidden + h * HD, md.y);
        tma_prefetch_2d(&maps.m[box - 1], 2 * hidden + h * HD, md.y);
      };
      for (int i = 0; i < 2 && i < my_items; ++i) { load_qk(i); load_v(i); }
      if (my_items > 2) prefetch_item(2);
      if (my_items > 3) prefetch_item(3);
      for (int i = 0; i + 2 < my_items; ++i) {
        // Q·Kᵀ(i) done: its Q and K tiles can be overwritten. The S buffer is single, its barrier is
        // not: Q·Kᵀ(i+1) needs nothing from this thread, so with one barrier a late look here could
        // find it two phases on, wait for Q·Kᵀ(i+2) — and that one needs the tiles loaded below.
        mbar_wait(bar(S_FULL, i & 1), (i >> 1) & 1);
        load_qk(i + 2);
        mbar_wait(bar(O_FULL, i & 1), (i >> 1) & 1);  // P·V(i) done: so can its V tile
        load_v(i + 2);
        if (i + 4 < my_items) prefetch_item(i + 4);
        stamp(i, 3);
      }
    }
  } else if (warp == 8) {
    // ===== MMA thread =====
    if (lane == 0) {
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
      // TMEM: S(i) in columns [0, 128); the softmax threads overwrite it in place with P(i) (packed
      // bf16: keys 0-63 in columns [0, 32), keys 64-127 in columns [64, 96)), which P·V(i) then takes
      // as its A operand straight from tensor memory. O(i) goes to columns [128 + 64·(i&1), +64) and
      // is read out one item later. The tensor core executes MMAs in issue order, so Q·Kᵀ(i+1),
      // issued right behind P·V(i), overwrites P(i) only after P·V(i) has consumed it.
      auto issue_s = [&](int i) {
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES;
        const int nk = ((meta[i].x + 31) >> 5) * 32;
        const uint32_t idesc_s = umma_idesc_bf16(128, nk, 0, 0);
        const uint64_t qd = umma_desc_sw128(st, 16, 1024);
        const uint64_t kd = umma_desc_sw128(st + TILE_BYTES, 16, 1024);
        mbar_wait(bar(QK_FULL, s), (i >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) umma_bf16(tmem_base, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        umma_commit(bar(S_FULL, s));
      };
      if (my_items > 0) issue_s(0);
      for (int i = 0; i < my_items; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        const int nk16 = (meta[i].x + 31) >> 5 << 1;  // K=16 steps of P·V: keys rounded up to 32, / 16
        // B = V: MN-major, one 128-byte line per key, 16 keys = 2048 bytes (>>4 = 128) per step
        const uint64_t vd0 = umma_desc_sw128(base + s * STAGE_BYTES + 2 * TILE_BYTES, 8192, 1024);
        mbar_wait(bar(V_FULL, s), ph);
        // every softmax thread arrives here after it has written its part of P(i) and, before that,
        // after its read of O(i-2): O buffer i&1 is free
        mbar_wait(bar(P_FULL, s), ph);
        stamp(i, 0);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < nk16)
            umma_bf16_ts(tmem_base + 128 + s * 64, tmem_base + (k >> 2) * 64 + (k & 3) * 8, vd0 + k * 128u,
                         idesc_o, k != 0);
        }
        umma_commit(bar(O_FULL, s));
        stamp(i, 2);
        if (i + 1 < my_items) issue_s(i + 1);
        stamp(i, 1);
      }
    }
  } else {
    const int r = threadIdx.x & 127;    // query row
    const int half = threadIdx.x >> 7;  // keys [64*half, +64), output columns [32*half, +32)
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    float* red_max = red;
    float* red_sum = red + 256;
    const uint32_t tS = tmem_base + lane_base;
    // O(i) is normalised and stored one item late, in the shadow of Q·Kᵀ(i+2): what it needs is kept
    // in registers across the iteration
    float inv_prev = 0.f;
    bool write_prev = false, live_prev = false;
    __nv_bfloat16* dst_prev = out;
    auto store_o = [&](int i, float inv, bool write_row, bool live, __nv_bfloat16* dst) {
      if (!live) return;  // this warp's 32 query rows lie wholly beyond the sequence: nothing to read or store
      mbar_wait(bar(O_FULL, i & 1), (i >> 1) & 1);
      if (threadIdx.x == 0) stamp(i, 6);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(tS + 128 + (i & 1) * 64 + half * 32, v);
      tmem_ld_wait();
      if (write_row) {
#pragma unroll
        for (int j = 0; j < 4; j += 2) {
          uint4 w[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int o = (j + u) * 8;
            w[u].x = pack_bf16x2(__uint_as_float(v[o + 0]) * inv, __uint_as_float(v[o + 1]) * inv);
            w[u].y = pack_bf16x2(__uint_as_float(v[o + 2]) * inv, __uint_as_float(v[o + 3]) * inv);
            w[u].z = pack_bf16x2(__uint_as_float(v[o + 4]) * inv, __uint_as_float(v[o + 5]) * inv);
            w[u].w = pack_bf16x2(__uint_as_float(v[o + 6]) * inv, __uint_as_float(v[o + 7]) * inv);
          }
          st_global_256(dst + j * 8, w[0], w[1]);  // 32-byte aligned: hidden and the head offset are multiples of 64
        }
      }
      if (threadIdx.x == 0) stamp(i, 7);
    };
    for (int i = 0; i < my_items; ++i) {
      const int s = i & 1;
      const int idx = blockIdx.x + i * gridDim.x;
      const int h = idx % heads;
      const int2 md = meta[i];
      const int len = md.x, row0 = md.y;
      // A warp owns query rows [32·(warp&3), +32). Rows at or beyond the sequence's TMA box hold stale
      // shared memory and are never stored, so a warp whose rows all lie there skips the TMEM reads,
      // the exponentials and the P write (its P rows stay whatever they were: P·V's output rows are
      // independent and those rows are not written either) and only keeps the barriers moving. At the
      // synthetic corpus's mean length of 70 tokens that is about a third of the softmax work.
      const bool live = (warp & 3) * 32 < len;
      const int kv = live ? min(64, max(0, len - 64 * half)) : 0;  // valid keys in this thread's half row
      mbar_wait(bar(S_FULL, s), (i >> 1) & 1);
      if (threadIdx.x == 0) stamp(i, 4);
      tc_fence_after();
      float mx = kv > 0 ? softmax_max64(tS + 64 * half, kv) : -INFINITY;
      red_max[half * 128 + r] = mx;
      softmax_bar();
      mx = fmaxf(mx, red_max[(half ^ 1) * 128 + r]);
      float rowsum = kv > 0 ? softmax_exp64_tmem(tS + 64 * half, kv, sl2, mx) : 0.f;
      red_sum[half * 128 + r] = rowsum;
      tc_fence_before();  // orders this thread's TMEM traffic (S(i) read, P(i) written, O(i-2) read)
      mbar_arrive(bar(P_FULL, s));
      if (threadIdx.x == 0) stamp(i, 5);
      softmax_bar();
      rowsum += red_sum[(half ^ 1) * 128 + r];
      // the previous item's output, while the tensor core works on S(i+1) and P·V(i)
      if (i > 0) store_o(i - 1, inv_prev, write_prev, live_prev, dst_prev);
      live_prev = live;
      inv_prev = rowsum > 0.f ? 1.0f / rowsum : 0.f;
      // rows past the sequence's length are pad (padded layout) or belong to the next sequence
      // (packed layout): never written
      write_prev = r < len;
      dst_prev = out + static_cast<size_t>(row0 + r) * hidden + h * HD + half * 32;
    }
    if (my_items > 0) store_o(my_items - 1, inv_prev, write_prev, live_prev, dst_prev);
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// ================================================================================================
// S = 256 .. 512: one ring step = (sequence, head, query block, key block)
// ================================================================================================
struct AttStep {
  int item, kb, nkb, h, qb, len, row0;
  bool valid;
};

__device__ __forceinline__ void att_step_next(AttStep& it, const int32_t* __restrict__ lens,
                                              const int32_t* __restrict__ cu, int S, int heads, int qblocks,
                                              int n_items) {
  if (++it.kb < it.nkb) return;
  for (;;) {
    it.item += gridDim.x;
    if (it.item >= n_items) { it.valid = false; return; }
    int t = it.item;
    it.qb = t % qblocks; t /= qblocks;
    it.h = t % heads;
    const int b = t / heads;
    int len = lens[b];
    len = len < 0 ? 0 : (len > S ? S : len);
    if (it.qb * QB >= len) continue;  // query block of pad rows only
    it.len = len;
    it.nkb = (len + KB - 1) / KB;
    it.kb = 0;
    it.row0 = cu ? cu[b] : b * S;
    return;
  }
}

__device__ __forceinline__ void att_step_init(AttStep& it, const int32_t* __restrict__ lens,
                                              const int32_t* __restrict__ cu, int S, int heads, int qblocks,
                                              int n_items) {
  it.item = static_cast<int>(blockIdx.x) - static_cast<int>(gridDim.x);
  it.kb = 0; it.nkb = 0; it.valid = true;
  att_step_next(it, lens, cu, S, heads, qblocks, n_items);
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_stream_kernel(const __grid_constant__ AttMaps maps, __nv_bfloat16* __restrict__ out,
                        const int32_t* __restrict__ lens, int S, int hidden, int heads, int qblocks,
                        int n_items, const int32_t* __restrict__ cu) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + OFF_BARS;
  auto bar = [&](int which, int s) { return bars + 8u * (which * 2 + s); };
  const uint32_t tmem_slot = bars + 8u * 10;
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_aligned + OFF_BARS + 80);
  float* red = reinterpret_cast<float*>(smem_aligned + OFF_RED);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == ATT_SOFTMAX_THREADS) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(QKV_FULL, s), 1);
      mbar_init(bar(S_FULL, s), 1);
      mbar_init(bar(P_FULL, s), ATT_SOFTMAX_THREADS);
      mbar_init(bar(O_FULL, s), 1);
      mbar_init(bar(FREE, s), ATT_SOFTMAX_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 8) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  auto kbox_of = [&](const AttStep& it) { return (min(KB, it.len - it.kb * KB) + 31) >> 5; };
  if (warp == 9) {
    // ===== TMA thread =====
    if (lane == 0) {
      auto load_step = [&](const AttStep& it, int i) {
        const int qbox = (min(QB, it.len - it.qb * QB) + 31) >> 5;
        const int kbox = kbox_of(it);
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES;
        mbar_expect_tx(bar(QKV_FULL, s), (qbox + 2 * kbox) * 32 * 128);
        tma_load_2d(st, &maps.m[qbox - 1], bar(QKV_FULL, s), it.h * HD, it.row0 + it.qb * QB);
        tma_load_2d(st + TILE_BYTES, &maps.m[kbox - 1], bar(QKV_FULL, s), hidden + it.h * HD,
                    it.row0 + it.kb * KB);
        tma_load_2d(st + 2 * TILE_BYTES, &maps.m[kbox - 1], bar(QKV_FULL, s), 2 * hidden + it.h * HD,
                    it.row0 + it.kb * KB);
      };
      AttStep ld;
      att_step_init(ld, lens, cu, S, heads, qblocks, n_items);
      if (ld.valid) { load_step(ld, 0); att_step_next(ld, lens, cu, S, heads, qblocks, n_items); }
      if (ld.valid) { load_step(ld, 1); att_step_next(ld, lens, cu, S, heads, qblocks, n_items); }
      for (int i = 0; ld.valid; ++i) {
        mbar_wait(bar(O_FULL, i & 1), (i >> 1) & 1);  // smem of the stage is free once P·V(i) has completed
        load_step(ld, i + 2);
        att_step_next(ld, lens, cu, S, heads, qblocks, n_items);
      }
    }
  } else if (warp == 8) {
    // ===== MMA thread =====
    if (lane == 0) {
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
      auto issue_s = [&](const AttStep& it, int i) {
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES;
        const uint32_t idesc_s = umma_idesc_bf16(128, kbox_of(it) * 32, 0, 0);
        const uint64_t qd = umma_desc_sw128(st, 16, 1024);
        const uint64_t kd = umma_desc_sw128(st + TILE_BYTES, 16, 1024);
        mbar_wait(bar(QKV_FULL, s), (i >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16(tmem_base + s * 128, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        umma_commit(bar(S_FULL, s));
      };
      AttStep si, pv;
      att_step_init(si, lens, cu, S, heads, qblocks, n_items);
      pv = si;
      if (si.valid) { issue_s(si, 0); att_step_next(si, lens, cu, S, heads, qblocks, n_items); }
      for (int i = 0; pv.valid; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        const uint32_t st = base + s * STAGE_BYTES;
        const int nk16 = kbox_of(pv) << 1;
        const uint64_t pd0 = umma_desc_sw128(st, 16, 1024);
        const uint64_t vd0 = umma_desc_sw128(st + 2 * TILE_BYTES, 8192, 1024);
        mbar_wait(bar(P_FULL, s), ph);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < nk16)
            umma_bf16(tmem_base + s * 128, pd0 + ((k >> 2) * (TILE_BYTES >> 4) + (k & 3) * 2), vd0 + k * 128u,
                      idesc_o, k != 0);
        }
        umma_commit(bar(O_FULL, s));
        if (si.valid) { issue_s(si, i + 1); att_step_next(si, lens, cu, S, heads, qblocks, n_items); }
        att_step_next(pv, lens, cu, S, heads, qblocks, n_items);
      }
    }
  } else {
    const int r = threadIdx.x & 127;
    const int half = threadIdx.x >> 7;
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    float* red_max = red;
    float* red_sum = red + 256;
    float m_run = -INFINITY, l_run = 0.f;
    float o[32];  // this thread's 32 output columns of its row, f32, rescaled block by block
    AttStep it;
    att_step_init(it, lens, cu, S, heads, qblocks, n_items);
    for (int i = 0; it.valid; ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      if (it.kb == 0) {
        m_run = -INFINITY; l_run = 0.f;
#pragma unroll
        for (int e = 0; e < 32; ++e) o[e] = 0.f;
      }
      const int kvalid = min(KB, it.len - it.kb * KB);
      const int kv = min(64, max(0, kvalid - 64 * half));
      const uint32_t tS = tmem_base + s * 128 + lane_base;
      uint8_t* tile_row = smem_aligned + s * STAGE_BYTES + half * TILE_BYTES + r * 128;
      mbar_wait(bar(S_FULL, s), ph);
      tc_fence_after();
      float mx = kv > 0 ? softmax_max64(tS + 64 * half, kv) : -INFINITY;
      red_max[half * 128 + r] = mx;
      softmax_bar();
      mx = fmaxf(mx, red_max[(half ^ 1) * 128 + r]);
      const float m_new = fmaxf(m_run, mx);
      float rowsum = kv > 0 ? softmax_exp64(tS + 64 * half, kv, sl2, m_new, tile_row, r) : 0.f;
      red_sum[half * 128 + r] = rowsum;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar(P_FULL, s));
      softmax_bar();
      rowsum += red_sum[(half ^ 1) * 128 + r];
      const float alpha = ex2_approx((m_run - m_new) * sl2);
      l_run = l_run * alpha + rowsum;
      m_run = m_new;

      mbar_wait(bar(O_FULL, s), ph);
      tc_fence_after();
      {
        uint32_t v[32];
        tmem_ld_32x32(tS + half * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = fmaf(o[j], alpha, __uint_as_float(v[j]));
      }
      tc_fence_before();
      mbar_arrive(bar(FREE, s));

      if (it.kb + 1 == it.nkb) {
        const int qvalid = min(QB, it.len - it.qb * QB);
        if (r < qvalid) {
          const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
          __nv_bfloat16* dst =
              out + static_cast<size_t>(it.row0 + it.qb * QB + r) * hidden + it.h * HD + half * 32;
#pragma unroll
          for (int j = 0; j < 4; j += 2) {
            uint4 w[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int c = (j + u) * 8;
              w[u].x = pack_bf16x2(o[c + 0] * inv, o[c + 1] * inv);
              w[u].y = pack_bf16x2(o[c + 2] * inv, o[c + 3] * inv);
              w[u].z = pack_bf16x2(o[c + 4] * inv, o[c + 5] * inv);
              w[u].w = pack_bf16x2(o[c + 6] * inv, o[c + 7] * inv);
            }
            st_global_256(dst + j * 8, w[0], w[1]);
          }
        }
      }
      att_step_next(it, lens, cu, S, heads, qblocks, n_items);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}


// ================================================================================================
// S = 256 .. 512, second design: one CTA per SM, two query blocks of one (sequence, head) in flight
// ================================================================================================
// The streaming kernel above is a serial chain per key block — Q·Kᵀ, softmax, P·V, then the softmax
// threads read O out of tensor memory to rescale it in registers before the next block may start —
// and runs at ~220 TFLOP/s. This kernel keeps the chain off the tensor core:
//   * a work item is (sequence, head, PAIR of 128-query blocks). Softmax group g (eight warps, two
//     threads per query row) owns query block g; the K and V tiles of a key block are loaded once
//     and used by both groups (half the shared-memory traffic per FLOP);
//   * the single MMA thread ping-pongs: when P_g(j) is ready it issues P_g(j)·V(j) and, right behind
//     it, Q_g·K(j+1)ᵀ, so group g's next scores are being computed while the other group is in its
//     exponentials; the tensor core never waits for a softmax as long as one group is ahead;
//   * P never touches shared memory (written as packed bf16 over the S columns, A operand of P·V
//     from tensor memory, as in the S = 128 kernel);
//   * O stays in tensor memory for the whole item: P·V accumulates across key blocks. The running
//     maximum is only replaced when a block's maximum exceeds it by more than 2^8 in the exponent
//     (then the softmax threads scale O and their partial denominators once, in place); otherwise
//     P is taken relative to the stale maximum, which is exact after the final division by the
//     denominator accumulated with the same reference and cannot overflow (P <= 2^8);
//   * Q tiles are double-buffered across items and K/V tiles run through a four-stage ring that the
//     TMA thread fills across item boundaries, so the next item's first Q·Kᵀ is issued right behind
//     the last P·V of the current one.
// Tensor memory (512 columns): S/P of group 0 at [0,128), of group 1 at [128,256), O at [256,320) and
// [320,384). Per (query pair, key block) the tensor core has 4 x 256 cycles of work and the MUFU
// 2 x 1,024 cycles of exponentials, so the exponentials bound this kernel at about half the tensor peak.
constexpr int PAIR_THREADS = 576;  // warps 0-7 softmax group 0, 8-15 group 1, 16 MMA issue, 17 TMA
constexpr int PAIR_STAGES = 4;     // K/V ring
constexpr int PAIR_MAX_ITEMS = 1024;
constexpr uint32_t PAIR_OFF_KV = 4 * TILE_BYTES;  // after Q[2 stages][2 groups]
constexpr uint32_t PAIR_OFF_BARS = PAIR_OFF_KV + PAIR_STAGES * 2 * TILE_BYTES;
constexpr uint32_t PAIR_OFF_RED = PAIR_OFF_BARS + 256;             // float [2 groups][2 halves][128], max then sum
constexpr uint32_t PAIR_OFF_META = PAIR_OFF_RED + 2 * 2 * 2 * 128 * 4;  // int2[PAIR_MAX_ITEMS]: (len, first row)
constexpr uint32_t PAIR_SMEM = 1024 + PAIR_OFF_META + PAIR_MAX_ITEMS * 8;
enum { PB_QFULL = 0, PB_QFREE = 2, PB_KFULL = 4, PB_VFULL = 8, PB_KVFREE = 12, PB_SFULL = 16, PB_PFULL = 18,
       PB_OFULL = 20, PB_TMEM = 22 };

struct PairStep {
  int i;       // index into this CTA's item table
  int n_item;  // how many items this CTA has started before this one
  int j, nkb, h, qp, len, row0;
  bool valid;
};

__device__ __forceinline__ void pair_next_item(PairStep& it, const int2* meta, int my_items, int heads, int qpairs) {
  for (;;) {
    if (++it.i >= my_items) { it.valid = false; return; }
    const int2 md = meta[it.i];
    int t = static_cast<int>(blockIdx.x) + it.i * static_cast<int>(gridDim.x);
    it.qp = t % qpairs; t /= qpairs;
    it.h = t % heads;
    if (it.qp * 2 * QB >= md.x) continue;  // both query blocks beyond the sequence
    it.len = md.x; it.row0 = md.y;
    it.nkb = (md.x + KB - 1) / KB;
    it.j = 0;
    ++it.n_item;
    return;
  }
}
__device__ __forceinline__ void pair_next(PairStep& it, const int2* meta, int my_items, int heads, int qpairs) {
  if (++it.j < it.nkb) return;
  pair_next_item(it, meta, my_items, heads, qpairs);
}
__device__ __forceinline__ void pair_init(PairStep& it, const int2* meta, int my_items, int heads, int qpairs) {
  it.i = -1; it.n_item = -1; it.j = 0; it.nkb = 0; it.valid = true;
  pair_next_item(it, meta, my_items, heads, qpairs);
}
// query rows of group g's block in this item (<= 0: the group has nothing to do in this item)
__device__ __forceinline__ int pair_qrows(const PairStep& it, int g) {
  return min(QB, it.len - (it.qp * 2 + g) * QB);
}

__global__ void __launch_bounds__(PAIR_THREADS, 1)
attention_pair_kernel(const __grid_constant__ AttMaps maps, __nv_bfloat16* __restrict__ out,
                      const int32_t* __restrict__ lens, int S, int hidden, int heads, int qpairs, int n_items,
                      const int32_t* __restrict__ cu) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + PAIR_OFF_BARS;
  auto bar = [&](int which, int s) { return bars + 8u * (which + s); };
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_aligned + PAIR_OFF_BARS + 8 * PB_TMEM);
  float* red = reinterpret_cast<float*>(smem_aligned + PAIR_OFF_RED);
  int2* meta = reinterpret_cast<int2*>(smem_aligned + PAIR_OFF_META);
  auto q_tile = [&](int qs, int g) { return base + static_cast<uint32_t>(qs * 2 + g) * TILE_BYTES; };
  auto k_tile = [&](int st) { return base + PAIR_OFF_KV + static_cast<uint32_t>(st) * 2 * TILE_BYTES; };
  auto v_tile = [&](int st) { return k_tile(st) + TILE_BYTES; };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int my_items = (n_items - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);
  for (int i = threadIdx.x; i < my_items; i += blockDim.x) {
    const int idx = blockIdx.x + i * gridDim.x;
    const int b = idx / (heads * qpairs);
    int len = lens[b];
    len = len < 0 ? 0 : (len > S ? S : len);
    meta[i] = make_int2(len, cu ? cu[b] : b * S);
  }
  if (threadIdx.x == 16 * 32) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(PB_QFULL, s), 1);
      mbar_init(bar(PB_QFREE, s), 1);
      mbar_init(bar(PB_SFULL, s), 1);
      mbar_init(bar(PB_PFULL, s), ATT_SOFTMAX_THREADS);
      mbar_init(bar(PB_OFULL, s), 1);
    }
    for (int s = 0; s < PAIR_STAGES; ++s) {
      mbar_init(bar(PB_KFULL, s), 1);
      mbar_init(bar(PB_VFULL, s), 1);
      mbar_init(bar(PB_KVFREE, s), 1);
    }
    fence_barrier_init();
  }
  if (warp == 16) tmem_alloc<512>(bars + 8u * PB_TMEM);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 17) {
    // ===== TMA thread: Q tiles per item, K / V tiles per step, running ahead through the ring =====
    if (lane == 0) {
      PairStep ld;
      pair_init(ld, meta, my_items, heads, qpairs);
      for (int t = 0; ld.valid; ++t) {
        const int st = t % PAIR_STAGES;
        const uint32_t ph = (t / PAIR_STAGES) & 1;
        if (ld.j == 0) {
          const int qs = ld.n_item & 1;
          mbar_wait(bar(PB_QFREE, qs), ((ld.n_item >> 1) & 1) ^ 1u);
          const int box0 = (pair_qrows(ld, 0) + 31) >> 5;
          const int rows1 = pair_qrows(ld, 1);
          const int box1 = rows1 > 0 ? (rows1 + 31) >> 5 : 0;
          mbar_expect_tx(bar(PB_QFULL, qs), (box0 + box1) * 32 * 128);
          tma_load_2d(q_tile(qs, 0), &maps.m[box0 - 1], bar(PB_QFULL, qs), ld.h * HD, ld.row0 + ld.qp * 2 * QB);
          if (box1)
            tma_load_2d(q_tile(qs, 1), &maps.m[box1 - 1], bar(PB_QFULL, qs), ld.h * HD,
                        ld.row0 + ld.qp * 2 * QB + QB);
        }
        mbar_wait(bar(PB_KVFREE, st), ph ^ 1u);
        const int kbox = (min(KB, ld.len - ld.j * KB) + 31) >> 5;
        mbar_expect_tx(bar(PB_KFULL, st), kbox * 32 * 128);
        tma_load_2d(k_tile(st), &maps.m[kbox - 1], bar(PB_KFULL, st), hidden + ld.h * HD, ld.row0 + ld.j * KB);
        mbar_expect_tx(bar(PB_VFULL, st), kbox * 32 * 128);
        tma_load_2d(v_tile(st), &maps.m[kbox - 1], bar(PB_VFULL, st), 2 * hidden + ld.h * HD, ld.row0 + ld.j * KB);
        pair_next(ld, meta, my_items, heads, qpairs);
      }
    }
  } else if (warp == 16) {
    // ===== MMA thread =====
    if (lane == 0) {
      constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
      auto kbox_of = [&](const PairStep& it) { return (min(KB, it.len - it.j * KB) + 31) >> 5; };
      // S_g(step) = Q_g · K(step)ᵀ into columns [128 g, +128)
      auto issue_qk = [&](const PairStep& it, int t, int g) {
        const int st = t % PAIR_STAGES;
        const int qs = it.n_item & 1;
        if (it.j == 0) mbar_wait(bar(PB_QFULL, qs), (it.n_item >> 1) & 1);
        mbar_wait(bar(PB_KFULL, st), (t / PAIR_STAGES) & 1);
        tc_fence_after();
        const uint32_t idesc_s = umma_idesc_bf16(128, kbox_of(it) * 32, 0, 0);
        const uint64_t qd = umma_desc_sw128(q_tile(qs, g), 16, 1024);
        const uint64_t kd = umma_desc_sw128(k_tile(st), 16, 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16(tmem_base + g * 128, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        umma_commit(bar(PB_SFULL, g));
      };
      PairStep cur, nxt;
      pair_init(cur, meta, my_items, heads, qpairs);
      nxt = cur;
      if (nxt.valid) pair_next(nxt, meta, my_items, heads, qpairs);
      uint32_t np[2] = {0, 0};  // P_FULL phases consumed per group
      if (cur.valid) {
        issue_qk(cur, 0, 0);
        if (pair_qrows(cur, 1) > 0) issue_qk(cur, 0, 1);
      }
      for (int t = 0; cur.valid; ++t) {
        const int st = t % PAIR_STAGES;
        const uint32_t ph = (t / PAIR_STAGES) & 1;
        const int last_g = pair_qrows(cur, 1) > 0 ? 1 : 0;
        const int nk16 = kbox_of(cur) << 1;
        const bool last_block = cur.j + 1 == cur.nkb;
        // B = V: MN-major, one 128-byte line per key, 16 keys = 2048 bytes (>>4 = 128) per step
        const uint64_t vd0 = umma_desc_sw128(v_tile(st), 8192, 1024);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          if (g <= last_g) {
            // every thread of group g arrives after it has written its part of P_g(step) — and, at the
            // first block of an item, after its read-out of the previous item's O_g
            mbar_wait(bar(PB_PFULL, g), np[g] & 1);
            ++np[g];
            mbar_wait(bar(PB_VFULL, st), ph);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if (k < nk16)
                umma_bf16_ts(tmem_base + 256 + g * 64, tmem_base + g * 128 + (k >> 2) * 64 + (k & 3) * 8,
                             vd0 + k * 128u, idesc_o, (cur.j > 0 || k != 0) ? 1u : 0u);
            }
            if (last_block) umma_commit(bar(PB_OFULL, g));
            if (g == last_g) {
              umma_commit(bar(PB_KVFREE, st));  // K(step) and V(step) have been consumed by both groups
              if (last_block) umma_commit(bar(PB_QFREE, cur.n_item & 1));
            }
          }
          // right behind it, the same group's scores for the next step (the tensor core executes in
          // issue order: S_g is overwritten only after P_g·V has consumed it)
          if (nxt.valid && pair_qrows(nxt, g) > 0) issue_qk(nxt, t + 1, g);
        }
        cur = nxt;
        if (nxt.valid) pair_next(nxt, meta, my_items, heads, qpairs);
      }
    }
  } else {
    // ===== softmax group g: query block g of every item =====
    const int g = warp >> 3;
    const int r = threadIdx.x & 127;          // query row within the block
    const int half = (threadIdx.x >> 7) & 1;  // keys [64*half, +64) of a key block, output columns [32*half, +32)
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    constexpr float kRescaleAt = 8.0f;                   // exponent (log2) by which a block maximum may exceed the reference
    float* red_max = red + g * 256;
    float* red_sum = red + 512 + g * 256;
    const uint32_t tS = tmem_base + lane_base + g * 128;
    const uint32_t tO = tmem_base + lane_base + 256 + g * 64 + half * 32;
    const int group_bar = 1 + g;
    auto group_sync = [&]() { asm volatile("bar.sync %0, 256;" ::"r"(group_bar) : "memory"); };
    uint32_t ns = 0, ni = 0;  // S_FULL / O_FULL phases consumed
    float m_ref = 0.f, l_part = 0.f;
    PairStep it;
    pair_init(it, meta, my_items, heads, qpairs);
    while (it.valid) {
      const int qrows = pair_qrows(it, g);
      if (qrows <= 0) {  // this group sits the item out
        pair_next_item(it, meta, my_items, heads, qpairs);
        continue;
      }
      const bool live = (warp & 3) * 32 < qrows;  // warps whose rows are all pad only keep the barriers moving
      const int kvalid = min(KB, it.len - it.j * KB);
      const int kv = live ? min(64, max(0, kvalid - 64 * half)) : 0;
      mbar_wait(bar(PB_SFULL, g), ns & 1);
      ++ns;
      tc_fence_after();
      float mx = kv > 0 ? softmax_max64(tS + 64 * half, kv) : -INFINITY;
      red_max[half * 128 + r] = mx;
      group_sync();
      mx = fmaxf(mx, red_max[(half ^ 1) * 128 + r]);
      if (live) {
        if (it.j == 0) {
          m_ref = mx;
          l_part = 0.f;
        } else if (__any_sync(0xffffffffu, (mx - m_ref) * sl2 > kRescaleAt)) {
          // S_FULL of this step implies P·V of the previous one has completed (commit covers every
          // earlier MMA), and P·V of this step waits for our arrival below: O is ours to scale.
          // The tensor-memory accesses are warp-wide, so the whole warp takes this path as soon as
          // one of its rows needs it; rows that do not keep their reference (alpha = 1).
          const bool grow = (mx - m_ref) * sl2 > kRescaleAt;
          const float alpha = grow ? ex2_approx((m_ref - mx) * sl2) : 1.0f;
          uint32_t v[32];
          tmem_ld_32x32(tO, v);
          tmem_ld_wait();
          uint32_t lo[16], hi[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            lo[e] = __float_as_uint(__uint_as_float(v[e]) * alpha);
            hi[e] = __float_as_uint(__uint_as_float(v[16 + e]) * alpha);
          }
          tmem_st_32x16(tO, lo);
          tmem_st_32x16(tO + 16, hi);
          tmem_st_wait();
          l_part *= alpha;
          if (grow) m_ref = mx;
        }
        if (kv > 0) l_part += softmax_exp64_tmem(tS + 64 * half, kv, sl2, m_ref);
      }
      tc_fence_before();
      mbar_arrive(bar(PB_PFULL, g));
      if (it.j + 1 == it.nkb) {
        // last key block: wait for the item's last P·V, normalise, store
        red_sum[half * 128 + r] = l_part;
        group_sync();
        const float l = l_part + red_sum[(half ^ 1) * 128 + r];
        mbar_wait(bar(PB_OFULL, g), ni & 1);
        ++ni;
        tc_fence_after();
        if (live) {
          uint32_t v[32];
          tmem_ld_32x32(tO, v);
          tmem_ld_wait();
          if (r < qrows) {
            const float inv = l > 0.f ? 1.0f / l : 0.f;
            __nv_bfloat16* dst = out + static_cast<size_t>(it.row0 + (it.qp * 2 + g) * QB + r) * hidden +
                                 it.h * HD + half * 32;
#pragma unroll
            for (int j = 0; j < 4; j += 2) {
              uint4 w[2];
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                const int c = (j + u) * 8;
                w[u].x = pack_bf16x2(__uint_as_float(v[c + 0]) * inv, __uint_as_float(v[c + 1]) * inv);
                w[u].y = pack_bf16x2(__uint_as_float(v[c + 2]) * inv, __uint_as_float(v[c + 3]) * inv);
                w[u].z = pack_bf16x2(__uint_as_float(v[c + 4]) * inv, __uint_as_float(v[c + 5]) * inv);
                w[u].w = pack_bf16x2(__uint_as_float(v[c + 6]) * inv, __uint_as_float(v[c + 7]) * inv);
              }
              st_global_256(dst + j * 8, w[0], w[1]);
            }
          }
        }
        tc_fence_before();  // the read of O is ordered before this thread's next P_FULL arrival
      }
      pair_next(it, meta, my_items, heads, qpairs);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 16) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace

// developer hook (dev_check): device buffer of 64 x 8 u64 that CTA 0 of the S=128 kernel fills
void attention_set_trace(unsigned long long* dev_buf) { g_att_trace = dev_buf; }

// `out` holds four tensor maps (box rows 32, 64, 96, 128) packed into AttentionMaps::raw.
int attention_tmaps(AttentionMaps* out, const void* qkv, int rows, int hidden) {
  static_assert(sizeof(AttentionMaps) == sizeof(AttMaps), "AttentionMaps layout");
  AttMaps* m = reinterpret_cast<AttMaps*>(out);
  for (int i = 0; i < 4; ++i) {
    int rc = make_tmap_bf16_2d(&m->m[i], qkv, rows, 3 * hidden, static_cast<uint64_t>(3 * hidden) * 2,
                               32 * (i + 1), 64);
    if (rc) return rc;
  }
  return B2_OK;
}

int attention_bf16(cudaStream_t stream, const AttentionMaps& tqkv, __nv_bfloat16* out,
                   const int32_t* lens, int B, int S, int hidden, const int32_t* cu) {
  if (B <= 0 || S <= 0 || (S % QB) != 0 || hidden <= 0 || (hidden % HD) != 0) {
    set_last_error("attention_bf16: needs S %% 128 == 0 and hidden %% 64 == 0 (got S=%d hidden=%d)", S,
                   hidden);
    return B2_ERR_ARG;
  }
  static bool configured[64] = {};
  int dev = 0;
  B2_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2_CUDA_OK(cudaFuncSetAttribute(attention_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM));
    B2_CUDA_OK(cudaFuncSetAttribute(attention_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM));
    B2_CUDA_OK(cudaFuncSetAttribute(attention_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    PAIR_SMEM));
    configured[dev] = true;
  }
  const int heads = hidden / HD, qblocks = S / QB;
  const long long items = 1ll * B * heads * qblocks;
  if (items > 0x7fffffffll) {
    set_last_error("attention_bf16: too many work items");
    return B2_ERR_ARG;
  }
  int sms = 0;
  B2_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const AttMaps& maps = reinterpret_cast<const AttMaps&>(tqkv);
  static const int per_sm = [] {  // developer knob: B2_ATT_CTAS_PER_SM=1 runs one CTA per SM
    const char* e = getenv("B2_ATT_CTAS_PER_SM");
    return e && atoi(e) == 1 ? 1 : 2;
  }();
  static const bool use_stream = [] {  // developer knob: B2_ATT_STREAM=1 keeps the first S > 128 kernel (A/B runs)
    const char* e = getenv("B2_ATT_STREAM");
    return e && atoi(e) == 1;
  }();
  int ctas = items < per_sm * sms ? static_cast<int>(items) : per_sm * sms;
  if (S == QB) {
    const int need = static_cast<int>((items + PERSIST_MAX_ITEMS - 1) / PERSIST_MAX_ITEMS);
    if (ctas < need) ctas = need;  // a CTA's item list must fit its shared-memory table
    attention_persist_kernel<<<ctas, ATT_THREADS, ATT_SMEM, stream>>>(maps, out, lens, S, hidden, heads,
                                                                      static_cast<int>(items), cu, g_att_trace);
  } else if (use_stream) {
    attention_stream_kernel<<<ctas, ATT_THREADS, ATT_SMEM, stream>>>(maps, out, lens, S, hidden, heads, qblocks,
                                                                     static_cast<int>(items), cu);
  } else {
    // one CTA per SM, work item = (sequence, head, pair of query blocks)
    const int qpairs = (qblocks + 1) / 2;
    const long long pitems = 1ll * B * heads * qpairs;
    int grid = pitems < sms ? static_cast<int>(pitems) : sms;
    const int need = static_cast<int>((pitems + PAIR_MAX_ITEMS - 1) / PAIR_MAX_ITEMS);
    if (grid < need) grid = need;  // a CTA's item list must fit its shared-memory table
    attention_pair_kernel<<<grid, PAIR_THREADS, PAIR_SMEM, stream>>>(maps, out, lens, S, hidden, heads, qpairs,
                                                                     static_cast<int>(pitems), cu);
  }
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int attention_set_trap_log(unsigned long long* mapped) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_trap_log, &mapped, sizeof(mapped)));
  return B2_OK;
}

}  // namespace b2
