// b2_attention.cu — masked self-attention for the encoder, head dim 64, on tcgen05 + TMA.
//
// One CTA handles one (sequence, head, 128-query block). For each 128-key block it runs
//   S = Q·Kᵀ            tcgen05.mma 128 x n x 64  (Q, K: K-major 128-byte-swizzled TMA tiles)
//   P = exp2((S-m)·c)    one thread per query row, S read from TMEM, keys >= len masked out
//   O_blk = P·V          tcgen05.mma 128 x 64 x n  (P: written to swizzled smem as the A operand,
//                                                  V: MN-major TMA tile, no transpose needed)
// with n = the block's valid keys rounded up to 32: the TMA boxes (four tensor maps with 32, 64, 96
// and 128 rows), the UMMA N of Q·Kᵀ and the K extent of P·V all shrink with the sequence, so a
// 40-token sentence moves and multiplies a third of what a 128-token one does. Rows of the tiles
// beyond the box hold stale shared memory; they are never read by an MMA whose result is kept.
//
// Two instantiations:
//   SINGLE (S <= 128): one key block. P overlays the Q|K tiles (dead once Q·Kᵀ has completed) and
//     O overlays S in TMEM, so a CTA needs 48 KiB of smem and 128 TMEM columns: 4 CTAs per SM
//     overlap each other's load -> MMA -> softmax -> MMA -> store chains.
//   MULTI (S = 256..512): online softmax over key blocks, running f32 output in registers; key
//     blocks and query blocks wholly beyond the sequence's length are skipped.
//
// No reference counterpart: the reference's model is an opaque TFLite graph
// (ops/map_classify_tpu.py:71-72). Semantics follow BertSelfAttention (scores/sqrt(d) + additive
// key mask, softmax in f32, context = probs·V).
#include <cstdlib>

#include "b2_kernels.h"

namespace b2 {

namespace {

constexpr int ATT_THREADS = 160;  // warps 0-3: one query row per thread; warp 4: TMA + MMA issue
constexpr int QB = 128;           // queries per CTA
constexpr int KB = 128;           // keys per block
constexpr int HD = 64;            // head dim
constexpr uint32_t TILE_BYTES = 128 * 64 * 2;  // 16 KiB: a [128 x 64] bf16 tile
constexpr uint32_t ATT_SMEM_SINGLE = 1024 + 3 * TILE_BYTES + 64;
constexpr uint32_t ATT_SMEM_MULTI = 1024 + 5 * TILE_BYTES + 64;

unsigned long long* g_att_trace = nullptr;

struct AttMaps {
  CUtensorMap m[4];  // box rows 32, 64, 96, 128 over the same [rows, 3H] buffer
};

template <bool SINGLE>
__global__ void __launch_bounds__(ATT_THREADS, SINGLE ? 4 : 2)
attention_kernel(const __grid_constant__ AttMaps maps, __nv_bfloat16* __restrict__ out,
                 const int32_t* __restrict__ lens, int S, int hidden, int heads, int qblocks,
                 const int32_t* __restrict__ cu) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sK = base + TILE_BYTES;
  const uint32_t sV = base + 2 * TILE_BYTES;
  // P: two [128 x 64] K-major tiles (keys 0-63, 64-127)
  constexpr uint32_t P_OFF = SINGLE ? 0u : 3 * TILE_BYTES;
  const uint32_t sP = base + P_OFF;
  const uint32_t bars = base + (SINGLE ? 3 : 5) * TILE_BYTES;
  const uint32_t q_full = bars, kv_full = bars + 8, s_full = bars + 16, p_full = bars + 24,
                 o_full = bars + 32, tmem_slot = bars + 40;
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  int cta = blockIdx.x;
  const int qb = cta % qblocks; cta /= qblocks;
  const int h = cta % heads;
  const int b = cta / heads;
  int len = lens[b];
  len = len < 0 ? 0 : (len > S ? S : len);
  const int nkb = (len + KB - 1) / KB;
  const int row0 = cu ? cu[b] : b * S;  // first token row of this sequence (packed or padded)
  if (qb * QB >= len) return;           // query block of pad rows only: nothing to compute or store
  const int qvalid = min(QB, len - qb * QB);
  constexpr uint32_t TMEM_COLS = SINGLE ? 128 : 256;

  if (threadIdx.x == 128) {
    mbar_init(q_full, 1);
    mbar_init(kv_full, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 128);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 4) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t tmem_S = tmem_base;                         // up to 128 columns
  const uint32_t tmem_O = tmem_base + (SINGLE ? 0u : 128u);  // 64 columns (SINGLE: over S)

  if (warp == 4) {
    if (lane == 0) {
      const int qbox = (qvalid + 31) >> 5;  // 1..4 -> 32..128 rows
      mbar_expect_tx(q_full, qbox * 32 * 128);
      tma_load_2d(sQ, &maps.m[qbox - 1], q_full, h * HD, row0 + qb * QB);
      for (int kb = 0; kb < nkb; ++kb) {
        const int kvalid = min(KB, len - kb * KB);
        const int kbox = (kvalid + 31) >> 5;
        const int nk = kbox * 32;  // keys this block's MMAs cover
        if (kb > 0) mbar_wait(o_full, (kb - 1) & 1);  // P·V of the previous block has read K/V/P
        mbar_expect_tx(kv_full, 2 * nk * 128);
        tma_load_2d(sK, &maps.m[kbox - 1], kv_full, hidden + h * HD, row0 + kb * KB);
        tma_load_2d(sV, &maps.m[kbox - 1], kv_full, 2 * hidden + h * HD, row0 + kb * KB);
        if (kb == 0) mbar_wait(q_full, 0);
        mbar_wait(kv_full, kb & 1);
        tc_fence_after();
        {
          const uint32_t idesc_s = umma_idesc_bf16(128, nk, 0, 0);
          const uint64_t qd = umma_desc_sw128(sQ, 16, 1024);
          const uint64_t kd = umma_desc_sw128(sK, 16, 1024);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) umma_bf16(tmem_S, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        }
        umma_commit(s_full);
        mbar_wait(p_full, kb & 1);
        tc_fence_after();
        constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
        for (int k = 0; k < nk / 16; ++k) {
          // A = P: K-major, 64 keys per 16 KiB tile, 32 bytes per K=16 step inside the swizzle atom
          const uint64_t pd = umma_desc_sw128(sP + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024);
          // B = V: MN-major, one 128-byte line per key, 16 keys = 2048 bytes per step
          const uint64_t vd = umma_desc_sw128(sV + k * 2048, 8192, 1024);
          umma_bf16(tmem_O, pd, vd, idesc_o, k != 0);
        }
        umma_commit(o_full);
      }
    }
  } else {
    // ===== softmax / output rows: thread r owns query row r =====
    const int r = threadIdx.x;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    float m_run = -INFINITY, l_run = 0.f;
    float o[SINGLE ? 1 : HD];
    if (!SINGLE) {
#pragma unroll
      for (int i = 0; i < (SINGLE ? 1 : HD); ++i) o[i] = 0.f;
    }

    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(s_full, kb & 1);
      tc_fence_after();
      const int kvalid = min(KB, len - kb * KB);
      float mx = -INFINITY;
      for (int c = 0; c < 4; ++c) {
        if (c * 32 < kvalid) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_S + lane_base + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c * 32 + j < kvalid) mx = fmaxf(mx, __uint_as_float(v[j]));
        }
      }
      // rows of stale shared memory (beyond the Q box) may hold anything, NaN included; their
      // results are never stored, they only must not disturb the arithmetic of this thread
      const float m_new = fmaxf(m_run, mx);
      const float alpha = exp2f((m_run - m_new) * sl2);
      float rowsum = 0.f;
      for (int c = 0; c < 4; ++c) {
        if (c * 32 < kvalid) {  // chunks at or beyond the key box are never read by the P·V MMA
          uint32_t packed[16];
          uint32_t v[32];
          tmem_ld_32x32(tmem_S + lane_base + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float p0 = (c * 32 + j < kvalid) ? exp2f((__uint_as_float(v[j]) - m_new) * sl2) : 0.f;
            float p1 = (c * 32 + j + 1 < kvalid) ? exp2f((__uint_as_float(v[j + 1]) - m_new) * sl2) : 0.f;
            rowsum += p0 + p1;
            packed[j >> 1] = pack_bf16x2(p0, p1);
          }
          // (SINGLE: P lands on the Q|K tiles; safe, s_full is only signalled once Q·Kᵀ completed)
          // keys [c*32, c*32+32) -> tile (c >> 1), 16-byte chunks ((c & 1) * 4 + j) ^ (r & 7)
          uint8_t* tile = smem_aligned + P_OFF + (c >> 1) * TILE_BYTES + r * 128;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int chunk = (((c & 1) * 4 + j) ^ (r & 7));
            *reinterpret_cast<uint4*>(tile + chunk * 16) =
                make_uint4(packed[j * 4], packed[j * 4 + 1], packed[j * 4 + 2], packed[j * 4 + 3]);
          }
        }
      }
      l_run = l_run * alpha + rowsum;
      m_run = m_new;
      fence_proxy_async_smem();  // P was written through the generic proxy; the MMA reads it async
      tc_fence_before();
      mbar_arrive(p_full);

      mbar_wait(o_full, kb & 1);
      tc_fence_after();
      if (!SINGLE) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_O + lane_base + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j)
            o[(SINGLE ? 0 : c * 32 + j)] = o[(SINGLE ? 0 : c * 32 + j)] * alpha + __uint_as_float(v[j]);
        }
      }
    }
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
    // rows past the sequence's length are pad (padded layout) or belong to the next sequence
    // (packed layout): never written
    __nv_bfloat16* dst = out + static_cast<size_t>(row0 + qb * QB + r) * hidden + h * HD;
    const bool store = r < qvalid;
    if (SINGLE) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_O + lane_base + c * 32, v);
        tmem_ld_wait();
        if (store) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(v[j * 8 + 0]) * inv, __uint_as_float(v[j * 8 + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(v[j * 8 + 2]) * inv, __uint_as_float(v[j * 8 + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(v[j * 8 + 4]) * inv, __uint_as_float(v[j * 8 + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(v[j * 8 + 6]) * inv, __uint_as_float(v[j * 8 + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + j * 8) = w;
          }
        }
      }
    } else if (store) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint4 w;
        w.x = pack_bf16x2(o[(SINGLE ? 0 : j * 8 + 0)] * inv, o[(SINGLE ? 0 : j * 8 + 1)] * inv);
        w.y = pack_bf16x2(o[(SINGLE ? 0 : j * 8 + 2)] * inv, o[(SINGLE ? 0 : j * 8 + 3)] * inv);
        w.z = pack_bf16x2(o[(SINGLE ? 0 : j * 8 + 4)] * inv, o[(SINGLE ? 0 : j * 8 + 5)] * inv);
        w.w = pack_bf16x2(o[(SINGLE ? 0 : j * 8 + 6)] * inv, o[(SINGLE ? 0 : j * 8 + 7)] * inv);
        *reinterpret_cast<uint4*>(dst + j * 8) = w;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// Softmax of one 128-key block for one query row, shared by the persistent kernels.
//   tS      TMEM address of this thread's row of S (lane bits included)
//   kvalid  keys of this block that belong to the sequence (1..128)
//   m_run   running row maximum from earlier key blocks (-inf for the first / only block)
//   prow    this row's 128 bytes in the first P tile (second tile at +TILE_BYTES), 128-byte swizzled
// Returns the new maximum and the sum of exp over this block; writes P (bf16) for every 32-key chunk
// that holds a valid key. TMEM loads are double-buffered: chunk c+1 is in flight while chunk c is
// reduced (tcgen05.wait::ld waits for everything issued, so the next load is issued right after it).
// A chunk that lies wholly inside the sequence takes the unmasked path: FMNMX3, packed FFMA2 / FADD2,
// two MUFU.EX2 and one bf16x2 convert per key pair; only the chunk that straddles `kvalid` is masked.
__device__ __forceinline__ void softmax_block(uint32_t tS, int kvalid, float sl2, float m_run, uint8_t* prow,
                                              int r, float& m_new, float& rowsum) {
  const int nch = (kvalid + 31) >> 5;
  uint32_t va[32], vb[32];
  float mx = -INFINITY;
  tmem_ld_32x32(tS, va);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (c < nch) {
      uint32_t(&cur)[32] = (c & 1) ? vb : va;
      uint32_t(&nxt)[32] = (c & 1) ? va : vb;
      tmem_ld_wait();
      if (c + 1 < nch) tmem_ld_32x32(tS + (c + 1) * 32, nxt);
      if (c * 32 + 32 <= kvalid) {
#pragma unroll
        for (int j = 0; j < 32; j += 2)
          mx = fmaxf(mx, fmaxf(__uint_as_float(cur[j]), __uint_as_float(cur[j + 1])));
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c * 32 + j < kvalid) mx = fmaxf(mx, __uint_as_float(cur[j]));
      }
    }
  }
  m_new = fmaxf(m_run, mx);
  const float moff = m_new * sl2;
  const uint64_t scale2 = f32x2(sl2, sl2), shift2 = f32x2(-moff, -moff);
  uint64_t sum2 = f32x2(0.f, 0.f);
  float sum1 = 0.f;
  tmem_ld_32x32(tS, va);  // second pass, chunk 0 (the one TMEM latency per block that stays exposed)
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (c < nch) {
      uint32_t(&cur)[32] = (c & 1) ? vb : va;
      uint32_t(&nxt)[32] = (c & 1) ? va : vb;
      uint32_t packed[16];
      tmem_ld_wait();
      if (c + 1 < nch) tmem_ld_32x32(tS + (c + 1) * 32, nxt);
      if (c * 32 + 32 <= kvalid) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float p0, p1;
          f32x2_split(ffma2(f32x2_bits(cur[j], cur[j + 1]), scale2, shift2), p0, p1);
          p0 = ex2_approx(p0);
          p1 = ex2_approx(p1);
          sum2 = fadd2(sum2, f32x2(p0, p1));
          packed[j >> 1] = pack_bf16x2(p0, p1);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float p0 = (c * 32 + j < kvalid) ? ex2_approx(fmaf(__uint_as_float(cur[j]), sl2, -moff)) : 0.f;
          const float p1 =
              (c * 32 + j + 1 < kvalid) ? ex2_approx(fmaf(__uint_as_float(cur[j + 1]), sl2, -moff)) : 0.f;
          sum1 += p0 + p1;
          packed[j >> 1] = pack_bf16x2(p0, p1);
        }
      }
      // keys [c*32, c*32+32) -> tile (c >> 1), 16-byte chunks ((c & 1) * 4 + j) ^ (r & 7)
      uint8_t* tile = prow + (c >> 1) * TILE_BYTES;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int chunk = (((c & 1) * 4 + j) ^ (r & 7));
        *reinterpret_cast<uint4*>(tile + chunk * 16) =
            make_uint4(packed[j * 4], packed[j * 4 + 1], packed[j * 4 + 2], packed[j * 4 + 3]);
      }
    }
  }
  float s0, s1;
  f32x2_split(sum2, s0, s1);
  rowsum = sum1 + s0 + s1;
}

// ---- persistent, software-pipelined variant for S = 128 (one key block per sequence) -------------
// Two CTAs per SM, each walking (sequence, head) items i = blockIdx.x, + gridDim.x, ... with two
// stages of {Q, K, V} tiles in shared memory and two S/O accumulators in TMEM:
//   control thread : loads for item i+2 are in flight while item i+1's Q·Kᵀ is issued right behind
//                    item i's P·V, so the tensor core never waits for a load and the softmax warps
//                    never wait for Q·Kᵀ;
//   softmax warps  : S_i (TMEM) -> P_i (smem, over the dead Q|K tiles of the same stage) ->
//                    O_i (TMEM, over S_i) -> scaled bf16 rows to global.
// Per-item set-up (TMEM allocation, barrier initialisation, CTA launch) is paid once per CTA
// instead of 12,288 times per layer.
constexpr uint32_t STAGE_BYTES_P = 3 * TILE_BYTES;
constexpr int PERSIST_MAX_ITEMS = 256;  // (sequence, head) items one CTA may walk; the launcher sizes the grid
constexpr uint32_t ATT_SMEM_PERSIST = 1024 + 2 * STAGE_BYTES_P + 128 + PERSIST_MAX_ITEMS * 8;

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_persist_kernel(const __grid_constant__ AttMaps maps, __nv_bfloat16* __restrict__ out,
                         const int32_t* __restrict__ lens, int S, int hidden, int heads, int n_items,
                         const int32_t* __restrict__ cu, unsigned long long* __restrict__ trace) {
  // developer timeline (dev_check "atttrace"): CTA 0 records %globaltimer at each pipeline event of
  // its first 64 items: [item][0..3] control thread (P seen, PV committed, O seen, load issued),
  // [item][4..7] softmax row 0 (S seen, P arrived, O seen, row stored)
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
  const uint32_t bars = base + 2 * STAGE_BYTES_P;
  // per stage s: qkv_full, s_full, p_full, o_full, free
  auto bar = [&](int which, int s) { return bars + 8u * (which * 2 + s); };
  enum { QKV_FULL = 0, S_FULL = 1, P_FULL = 2, O_FULL = 3, FREE = 4 };
  const uint32_t tmem_slot = bars + 8u * 10;
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int my_items = (n_items - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);
  // Per-item (length, first row) looked up ONCE into shared memory: the control thread's loop is a
  // chain of dependent steps, and a global load of lens[] / cu[] in it costs ~0.5 us each time
  // (measured with the dev_check timeline: 2 us of a 4.8 us item were such loads).
  int2* meta = reinterpret_cast<int2*>(smem_aligned + 2 * STAGE_BYTES_P + 128);
  for (int i = threadIdx.x; i < my_items; i += blockDim.x) {
    const int idx = blockIdx.x + i * gridDim.x;
    const int b = idx / heads;
    int len = lens[b];
    len = len < 1 ? 1 : (len > S ? S : len);
    meta[i] = make_int2(len, cu ? cu[b] : b * S);
  }

  if (threadIdx.x == 128) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(QKV_FULL, s), 1);
      mbar_init(bar(S_FULL, s), 1);
      mbar_init(bar(P_FULL, s), 128);
      mbar_init(bar(O_FULL, s), 1);
      mbar_init(bar(FREE, s), 128);
    }
    fence_barrier_init();
  }
  if (warp == 4) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 4) {
    if (lane == 0) {
      auto load_item = [&](int i) {
        const int idx = blockIdx.x + i * gridDim.x;
        const int h = idx % heads;
        const int2 md = meta[i];
        const int row0 = md.y;
        const int box = (md.x + 31) >> 5;
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        mbar_expect_tx(bar(QKV_FULL, s), 3 * box * 32 * 128);
        tma_load_2d(st, &maps.m[box - 1], bar(QKV_FULL, s), h * HD, row0);
        tma_load_2d(st + TILE_BYTES, &maps.m[box - 1], bar(QKV_FULL, s), hidden + h * HD, row0);
        tma_load_2d(st + 2 * TILE_BYTES, &maps.m[box - 1], bar(QKV_FULL, s), 2 * hidden + h * HD, row0);
      };
      // pull item i's tiles from DRAM into L2 ahead of its TMA load (no shared memory needed)
      auto prefetch_item = [&](int i) {
        const int idx = blockIdx.x + i * gridDim.x;
        const int h = idx % heads;
        const int2 md = meta[i];
        const int row0 = md.y;
        const int box = (md.x + 31) >> 5;
        tma_prefetch_2d(&maps.m[box - 1], h * HD, row0);
        tma_prefetch_2d(&maps.m[box - 1], hidden + h * HD, row0);
        tma_prefetch_2d(&maps.m[box - 1], 2 * hidden + h * HD, row0);
      };
      auto nk_of = [&](int i) { return ((meta[i].x + 31) >> 5) * 32; };
      auto issue_s = [&](int i) {
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        mbar_wait(bar(QKV_FULL, s), (i >> 1) & 1);
        tc_fence_after();
        const uint32_t idesc_s = umma_idesc_bf16(128, nk_of(i), 0, 0);
        const uint64_t qd = umma_desc_sw128(st, 16, 1024);
        const uint64_t kd = umma_desc_sw128(st + TILE_BYTES, 16, 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16(tmem_base + s * 128, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        umma_commit(bar(S_FULL, s));
      };
      if (my_items > 0) load_item(0);
      if (my_items > 1) load_item(1);
      if (my_items > 2) prefetch_item(2);
      if (my_items > 3) prefetch_item(3);
      if (my_items > 0) issue_s(0);
      for (int i = 0; i < my_items; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        const int nk = nk_of(i);
        mbar_wait(bar(P_FULL, s), ph);
        stamp(i, 0);
        tc_fence_after();
        constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
        for (int k = 0; k < nk / 16; ++k) {
          const uint64_t pd = umma_desc_sw128(st + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024);
          const uint64_t vd = umma_desc_sw128(st + 2 * TILE_BYTES + k * 2048, 8192, 1024);
          umma_bf16(tmem_base + s * 128, pd, vd, idesc_o, k != 0);
        }
        umma_commit(bar(O_FULL, s));
        if (i + 1 < my_items) issue_s(i + 1);  // right behind P·V: the softmax warps find S ready
        stamp(i, 1);
        if (i + 2 < my_items) {
          // stage s's shared memory is free as soon as P·V(i) has completed; its TMEM columns are
          // only needed again by S(i+2), which is issued one iteration later, after the softmax
          // threads have read O(i) (they arrive on P_FULL(i+1) only after that read)
          mbar_wait(bar(O_FULL, s), ph);
          stamp(i, 2);
          load_item(i + 2);
          if (i + 4 < my_items) prefetch_item(i + 4);
          stamp(i, 3);
        }
      }
    }
  } else {
    const int r = threadIdx.x;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    for (int i = 0; i < my_items; ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      const int idx = blockIdx.x + i * gridDim.x;
      const int h = idx % heads;
      const int2 md = meta[i];
      const int len = md.x, row0 = md.y;
      const uint32_t tS = tmem_base + s * 128 + lane_base;
      mbar_wait(bar(S_FULL, s), ph);
      if (r == 0) stamp(i, 4);
      tc_fence_after();
      float m_new, rowsum;
      softmax_block(tS, len, sl2, -INFINITY, smem_aligned + s * STAGE_BYTES_P + r * 128, r, m_new, rowsum);
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar(P_FULL, s));
      if (r == 0) stamp(i, 5);

      const float inv = rowsum > 0.f ? 1.0f / rowsum : 0.f;
      __nv_bfloat16* dst = out + static_cast<size_t>(row0 + r) * hidden + h * HD;
      const bool store = r < len;
      mbar_wait(bar(O_FULL, s), ph);
      if (r == 0) stamp(i, 6);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tS + c * 32, v);
        tmem_ld_wait();
        if (store) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 w;
            w.x = pack_bf16x2(__uint_as_float(v[j * 8 + 0]) * inv, __uint_as_float(v[j * 8 + 1]) * inv);
            w.y = pack_bf16x2(__uint_as_float(v[j * 8 + 2]) * inv, __uint_as_float(v[j * 8 + 3]) * inv);
            w.z = pack_bf16x2(__uint_as_float(v[j * 8 + 4]) * inv, __uint_as_float(v[j * 8 + 5]) * inv);
            w.w = pack_bf16x2(__uint_as_float(v[j * 8 + 6]) * inv, __uint_as_float(v[j * 8 + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + c * 32 + j * 8) = w;
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar(FREE, s));
      if (r == 0) stamp(i, 7);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

// ---- persistent variant for S = 256 .. 512 (several key blocks per query block) ------------------
// Same ring as attention_persist_kernel, but a ring step is one (sequence, head, query block, key
// block): Q is re-fetched with every key block (an L2 hit; it lets P overlay the Q|K tiles exactly as
// in the one-block kernel), and the softmax threads carry the running maximum, denominator and the
// f32 output row across the key blocks of a query block (online softmax), writing the row once.
struct AttStep {
  int item, kb, nkb, b, h, qb, len, row0;
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
    it.b = t / heads;
    int len = lens[it.b];
    len = len < 0 ? 0 : (len > S ? S : len);
    if (it.qb * QB >= len) continue;  // query block of pad rows only
    it.len = len;
    it.nkb = (len + KB - 1) / KB;
    it.kb = 0;
    it.row0 = cu ? cu[it.b] : it.b * S;
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
  const uint32_t bars = base + 2 * STAGE_BYTES_P;
  auto bar = [&](int which, int s) { return bars + 8u * (which * 2 + s); };
  enum { QKV_FULL = 0, S_FULL = 1, P_FULL = 2, O_FULL = 3, FREE = 4 };
  const uint32_t tmem_slot = bars + 8u * 10;
  uint8_t* smem_aligned = smem_raw + (base - raw);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 128) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(QKV_FULL, s), 1);
      mbar_init(bar(S_FULL, s), 1);
      mbar_init(bar(P_FULL, s), 128);
      mbar_init(bar(O_FULL, s), 1);
      mbar_init(bar(FREE, s), 128);
    }
    fence_barrier_init();
  }
  if (warp == 4) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 4) {
    if (lane == 0) {
      auto kbox_of = [&](const AttStep& it) { return (min(KB, it.len - it.kb * KB) + 31) >> 5; };
      auto load_step = [&](const AttStep& it, int i) {
        const int qbox = (min(QB, it.len - it.qb * QB) + 31) >> 5;
        const int kbox = kbox_of(it);
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        mbar_expect_tx(bar(QKV_FULL, s), (qbox + 2 * kbox) * 32 * 128);
        tma_load_2d(st, &maps.m[qbox - 1], bar(QKV_FULL, s), it.h * HD, it.row0 + it.qb * QB);
        tma_load_2d(st + TILE_BYTES, &maps.m[kbox - 1], bar(QKV_FULL, s), hidden + it.h * HD,
                    it.row0 + it.kb * KB);
        tma_load_2d(st + 2 * TILE_BYTES, &maps.m[kbox - 1], bar(QKV_FULL, s), 2 * hidden + it.h * HD,
                    it.row0 + it.kb * KB);
      };
      auto issue_s = [&](const AttStep& it, int i) {
        const int s = i & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        mbar_wait(bar(QKV_FULL, s), (i >> 1) & 1);
        tc_fence_after();
        const uint32_t idesc_s = umma_idesc_bf16(128, kbox_of(it) * 32, 0, 0);
        const uint64_t qd = umma_desc_sw128(st, 16, 1024);
        const uint64_t kd = umma_desc_sw128(st + TILE_BYTES, 16, 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16(tmem_base + s * 128, qd + 2u * k, kd + 2u * k, idesc_s, k != 0);
        umma_commit(bar(S_FULL, s));
      };
      AttStep ld, si, pv;
      att_step_init(ld, lens, cu, S, heads, qblocks, n_items);
      si = ld; pv = ld;
      if (ld.valid) { load_step(ld, 0); att_step_next(ld, lens, cu, S, heads, qblocks, n_items); }
      if (ld.valid) { load_step(ld, 1); att_step_next(ld, lens, cu, S, heads, qblocks, n_items); }
      if (si.valid) { issue_s(si, 0); att_step_next(si, lens, cu, S, heads, qblocks, n_items); }
      for (int i = 0; pv.valid; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        const uint32_t st = base + s * STAGE_BYTES_P;
        const int nk = kbox_of(pv) * 32;
        mbar_wait(bar(P_FULL, s), ph);
        tc_fence_after();
        constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);  // B (= V) is MN-major
        for (int k = 0; k < nk / 16; ++k) {
          const uint64_t pd = umma_desc_sw128(st + (k >> 2) * TILE_BYTES + (k & 3) * 32, 16, 1024);
          const uint64_t vd = umma_desc_sw128(st + 2 * TILE_BYTES + k * 2048, 8192, 1024);
          umma_bf16(tmem_base + s * 128, pd, vd, idesc_o, k != 0);
        }
        umma_commit(bar(O_FULL, s));
        if (si.valid) { issue_s(si, i + 1); att_step_next(si, lens, cu, S, heads, qblocks, n_items); }
        if (ld.valid) {
          mbar_wait(bar(FREE, s), ph);
          load_step(ld, i + 2);
          att_step_next(ld, lens, cu, S, heads, qblocks, n_items);
        }
        att_step_next(pv, lens, cu, S, heads, qblocks, n_items);
      }
    }
  } else {
    const int r = threadIdx.x;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    const float sl2 = 0.125f * 1.44269504088896340736f;  // 1/sqrt(64) * log2(e)
    float m_run = -INFINITY, l_run = 0.f;
    float o[HD];
    AttStep it;
    att_step_init(it, lens, cu, S, heads, qblocks, n_items);
    for (int i = 0; it.valid; ++i) {
      const int s = i & 1;
      const uint32_t ph = (i >> 1) & 1;
      if (it.kb == 0) {
        m_run = -INFINITY; l_run = 0.f;
#pragma unroll
        for (int e = 0; e < HD; ++e) o[e] = 0.f;
      }
      const int kvalid = min(KB, it.len - it.kb * KB);
      const uint32_t tS = tmem_base + s * 128 + lane_base;
      mbar_wait(bar(S_FULL, s), ph);
      tc_fence_after();
      float m_new, rowsum;
      softmax_block(tS, kvalid, sl2, m_run, smem_aligned + s * STAGE_BYTES_P + r * 128, r, m_new, rowsum);
      const float alpha = ex2_approx((m_run - m_new) * sl2);
      l_run = l_run * alpha + rowsum;
      m_run = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar(P_FULL, s));

      mbar_wait(bar(O_FULL, s), ph);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tS + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) o[c * 32 + j] = fmaf(o[c * 32 + j], alpha, __uint_as_float(v[j]));
      }
      tc_fence_before();
      mbar_arrive(bar(FREE, s));

      if (it.kb + 1 == it.nkb) {
        const int qvalid = min(QB, it.len - it.qb * QB);
        if (r < qvalid) {
          const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
          __nv_bfloat16* dst = out + static_cast<size_t>(it.row0 + it.qb * QB + r) * hidden + it.h * HD;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 w;
            w.x = pack_bf16x2(o[j * 8 + 0] * inv, o[j * 8 + 1] * inv);
            w.y = pack_bf16x2(o[j * 8 + 2] * inv, o[j * 8 + 3] * inv);
            w.z = pack_bf16x2(o[j * 8 + 4] * inv, o[j * 8 + 5] * inv);
            w.w = pack_bf16x2(o[j * 8 + 6] * inv, o[j * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(dst + j * 8) = w;
          }
        }
      }
      att_step_next(it, lens, cu, S, heads, qblocks, n_items);
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
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
    B2_CUDA_OK(cudaFuncSetAttribute(attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM_SINGLE));
    B2_CUDA_OK(cudaFuncSetAttribute(attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM_MULTI));
    B2_CUDA_OK(cudaFuncSetAttribute(attention_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM_PERSIST));
    B2_CUDA_OK(cudaFuncSetAttribute(attention_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    ATT_SMEM_PERSIST));
    configured[dev] = true;
  }
  const int heads = hidden / HD, qblocks = S / QB;
  const long long grid = 1ll * B * heads * qblocks;
  if (grid > 0x7fffffffll) {
    set_last_error("attention_bf16: grid too large");
    return B2_ERR_ARG;
  }
  const AttMaps& maps = reinterpret_cast<const AttMaps&>(tqkv);
  static const bool one_shot = [] { const char* e = getenv("B2_ATTN_ONESHOT"); return e && e[0] == '1'; }();
  if (S == QB && !one_shot) {
    int sms = 0;
    B2_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int items = B * heads;
    int ctas = items < 2 * sms ? items : 2 * sms;
    const int need = (items + PERSIST_MAX_ITEMS - 1) / PERSIST_MAX_ITEMS;  // keep a CTA's item list in smem
    if (ctas < need) ctas = need;
    attention_persist_kernel<<<ctas, ATT_THREADS, ATT_SMEM_PERSIST, stream>>>(maps, out, lens, S, hidden,
                                                                             heads, items, cu, g_att_trace);
  } else if (!one_shot) {
    int sms = 0;
    B2_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int items = static_cast<int>(grid);
    const int ctas = items < 2 * sms ? items : 2 * sms;
    attention_stream_kernel<<<ctas, ATT_THREADS, ATT_SMEM_PERSIST, stream>>>(maps, out, lens, S, hidden, heads,
                                                                            qblocks, items, cu);
  } else if (S == QB)
    attention_kernel<true><<<static_cast<unsigned>(grid), ATT_THREADS, ATT_SMEM_SINGLE, stream>>>(
        maps, out, lens, S, hidden, heads, qblocks, cu);
  else
    attention_kernel<false><<<static_cast<unsigned>(grid), ATT_THREADS, ATT_SMEM_MULTI, stream>>>(
        maps, out, lens, S, hidden, heads, qblocks, cu);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace b2
