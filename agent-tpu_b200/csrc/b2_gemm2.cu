// b2_gemm2.cu — the encoder GEMM on CTA pairs: tcgen05.mma.cta_group::2, 256 x 256 pair tile.
//
//   C[M,N] (bf16) = epilogue( A[M,K] · W[N,K]^T + bias[N] )      (same contract as b2_gemm.cu)
//
// Why a second kernel: with one CTA per 128 x 256 tile every SM pulls (128 + 256) x 64 bf16 = 48 KiB
// from L2 per 512 tensor-pipe cycles (96 B/clk/SM, 14 KB/clk for the chip) and the 4-stage ring only
// covers ~2,000 cycles of latency; ncu showed the tensor pipe 51% active with L2 and L1 both ~63%
// busy (profiles/r01_ncu_gemm.md). A CTA pair shares the operands of a 256 x 256 tile: each CTA
// loads its 128 rows of A and its 128 rows of W (32 KiB per K step, 64 B/clk/SM), the tensor cores
// of both SMs read both halves of W, and six 32 KiB stages fit where four 48 KiB ones did.
//
// Layout of the work
//   cluster (2,1,1); pair p = blockIdx.x / 2 walks tiles p, p + P, ... (P pairs = SMs / 2)
//   CTA rank r owns rows [m0 + 128 r, +128) of the tile: loads them (A) and W rows [n0 + 128 r, +128)
//   warp 16  TMA producer (both CTAs; bytes are counted on the LEADER's full barrier)
//   warp 17  MMA issuer (leader only): UMMA 256 x 256 x 16, accumulator rows 0-127 in the leader's
//            TMEM, 128-255 in the peer's; tcgen05.commit multicasts to both CTAs' barriers
//   warp 18  TMEM allocator (cta_group::2 alloc / dealloc, both CTAs)
//   warps 0-15 epilogue, each CTA drains its own 128 x 256 accumulator half; accumulator stages are
//            released on the leader's tmem-empty barrier (the peer arrives remotely)
//
// Epilogue data path. One thread owns one accumulator row (that is how tcgen05.ld hands TMEM out),
// so direct global stores are 32 rows x 16 bytes per warp instruction: 32 L1 wavefronts each, and
// ncu showed L1/TEX 60-64% busy with the K = 768 GEMMs stuck at 52-58% tensor-pipe activity while
// the K = 3072 GEMM, which hides the same epilogue behind a 4x longer main loop, ran at 84-90%
// (profiles/r01_ncu_gemm2.md). So the tile goes TMEM -> registers -> a 128-byte-swizzled shared
// staging block (conflict-free 16-byte writes) -> one TMA store per 128 x 64 block, and the
// residual block arrives the same way by TMA load into the staging block before it is needed.
#include "b2_kernels.h"

namespace b2 {

namespace {

constexpr int BM = 128;   // rows per CTA (pair tile: 256)
constexpr int BN = 256;   // columns of the pair tile
constexpr int BNH = 128;  // W rows each CTA loads
constexpr int BK = 64;
constexpr int STAGES = 5;
constexpr int THREADS = 640;
constexpr int EPI_WARPS = 16;
// Warp roles. The SM sub-partition arbiter prefers the highest warp id among eligible warps, so the
// single-thread roles that pace the tensor pipe (TMA producer, MMA issuer) sit ABOVE the sixteen
// epilogue warps: placed below them they were starved of issue slots whenever the epilogue was busy,
// and every instruction added to the epilogue slowed the main loop down.
constexpr int W_TMA = 16, W_MMA = 17, W_ALLOC = 18;
constexpr uint32_t A_BYTES = BM * BK * 2;   // 16 KiB
constexpr uint32_t B_BYTES = BNH * BK * 2;  // 16 KiB
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t OUT_BYTES = 128 * 64 * 2;  // one epilogue staging block: 128 rows x 64 bf16
constexpr uint32_t BAR_BYTES = 8 * (2 * STAGES + 4 + 4) + 16;
constexpr uint32_t SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + 4 * OUT_BYTES + BAR_BYTES;
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // shared::cluster address -> same offset in the even CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                              uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this smem offset in every CTA of `mask` once the MMAs issued so far are done
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"(mask)
      : "memory");
}
// TMA load whose completion bytes are counted on the leader CTA's barrier
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst_smem, const CUtensorMap* tmap, uint32_t bar,
                                                int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar & PEER_MASK), "r"(c0), "r"(c1)
      : "memory");
}
// arrive on the barrier at the same smem offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t"
      "}"
      ::"r"(bar), "r"(rank)
      : "memory");
}

template <int EPI>
__device__ __forceinline__ float epi_act(float v) {
  if (EPI == GEMM_EPI_BIAS_GELU || EPI == GEMM_EPI_LNIN_GELU) return gelu_erf(v);
  if (EPI == GEMM_EPI_BIAS_TANH) return tanhf(v);
  return v;
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm2_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmap_a,
                     const __grid_constant__ CUtensorMap tmap_b,
                     const __grid_constant__ CUtensorMap tmap_c,
                     const __grid_constant__ CUtensorMap tmap_r, const float* __restrict__ bias, int M,
                     int N, int K, const int32_t* __restrict__ m_dev, const GemmLnArgs ln) {
  extern __shared__ uint8_t smem_raw[];
  constexpr bool kLnIn = EPI == GEMM_EPI_LNIN || EPI == GEMM_EPI_LNIN_GELU;
  constexpr bool kLnRes = EPI == GEMM_EPI_LNRES_STATS;
  constexpr bool kResidual = EPI == GEMM_EPI_BIAS_RESIDUAL || kLnRes;
  if (m_dev) M = min(M, __ldg(m_dev));
  // developer timeline: pair 0's leader records %globaltimer per tile — [0] accumulator stage free,
  // [1] main loop issued (MMA thread); [2] staging block free, [3] accumulator ready, [4] residual
  // landed, [5] first half drained, [6] second half drained, [7] TMA store issued (epilogue warp 0)
  auto stamp = [&](int tile_it, int slot) {
    if (ln.trace && blockIdx.x == 0 && tile_it < 32) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      ln.trace[tile_it * 8 + slot] = t;
      ln.trace[256 + tile_it * 8 + slot] = static_cast<unsigned long long>(clock64());  // SM cycles
    }
  };
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t sA = base;
  const uint32_t sB = base + STAGES * A_BYTES;
  const uint32_t sOut = base + STAGES * STAGE_BYTES;  // four 16 KiB staging blocks, one per quarter
  const uint32_t bars = sOut + 4 * OUT_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + 2 + a); };
  auto rfull_bar = [&](int qt) { return bars + 8u * (2 * STAGES + 4 + qt); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 8);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  const int m_tiles = (M + 2 * BM - 1) / (2 * BM);
  const int n_tiles = (N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_c);
    if (EPI == GEMM_EPI_BIAS_RESIDUAL || EPI == GEMM_EPI_LNRES_STATS) tma_prefetch_desc(&tmap_r);
  }
  if (warp == W_MMA && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 2);   // leader's arrive.expect_tx + the peer producer's remote arrive
      mbar_init(empty_bar(s), 1);  // one multicast tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 2 * EPI_WARPS);  // epilogue warps of both CTAs (leader's copy is used)
    }
    for (int qt = 0; qt < 4; ++qt) mbar_init(rfull_bar(qt), 1);  // residual block landed
    fence_barrier_init();
  }
  cluster_sync_all();  // both CTAs' barriers exist before anything arrives on them remotely
  if (warp == W_ALLOC) tmem_alloc_2sm(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == W_TMA) {
    // ===== TMA producer (one thread in each CTA) =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += num_pairs) {
        const int m_blk = t / n_tiles, n_blk = t % n_tiles;
        const int arow = m_blk * 2 * BM + rank * BM;
        const int brow = n_blk * BN + rank * BNH;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          if (ln.debug_mode == 1 && (t != pair || kb >= STAGES)) {  // dev_check: MMA-only pace
            if (leader) mbar_arrive(full_bar(stage));
            else mbar_arrive_cluster(full_bar(stage), 0);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            continue;
          }
          if (leader) mbar_expect_tx(full_bar(stage), 2 * STAGE_BYTES);
          else mbar_arrive_cluster(full_bar(stage), 0);
          tma_load_2d_2sm(sA + stage * A_BYTES, &tmap_a, full_bar(stage), kb * BK, arow);
          tma_load_2d_2sm(sB + stage * B_BYTES, &tmap_b, full_bar(stage), kb * BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer: one thread of the leader CTA drives both SMs' tensor cores =====
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        stamp(it, 0);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_sw128(sA + stage * A_BYTES, 16, 1024);
          const uint64_t b_desc = umma_desc_sw128(sB + stage * B_BYTES, 16, 1024);
          if (ln.debug_mode != 2) {  // 2 (dev_check): TMA-only pace
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_2sm(d_tmem, a_desc + 2u * k, b_desc + 2u * k, idesc, (kb | k) != 0);
          }
          umma_commit_2sm(empty_bar(stage), 0b11);  // both CTAs' smem slots are free again
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit_2sm(tfull_bar(as), 0b11);  // both accumulator halves complete
        stamp(it, 1);
      }
    }
  } else if (warp < EPI_WARPS) {
    // ===== epilogue: each CTA drains its own 128 x 256 half, one 128 x 64 block per quarter group =====
    const int q = warp & 3;        // TMEM lane quadrant of this warp
    const int quarter = warp >> 2;  // 64-column quarter: warps 4*quarter .. 4*quarter+3 share a block
    const int r_local = q * 32 + lane;
    const bool elected = (q == 0) && (lane == 0);
    const uint32_t buf = sOut + quarter * OUT_BYTES;
    uint8_t* buf_ptr = smem_raw + (buf - raw) + r_local * 128;
    const uint32_t sw = static_cast<uint32_t>(r_local & 7);
    // Phase of this quarter's residual barrier. It flips with every tile in which the quarter loads a
    // residual block — not with every tile: when N is not a multiple of 256 (hidden 384, 640, 896) the
    // quarters beyond N sit out the last column tile of each row block.
    uint32_t rphase = 0;
    int it = 0;
    for (int t = pair; t < num_tiles; t += num_pairs, ++it) {
      const int m_blk = t / n_tiles, n_blk = t % n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const int row_base = m_blk * 2 * BM + rank * BM;
      const int row = row_base + r_local;
      const bool row_ok = row < M;
      const int colq = n_blk * BN + quarter * 64;
      const bool active = colq < N;  // quarters wholly beyond N only keep the barriers moving
      // the previous tile's TMA store must have finished reading this staging block
      if (elected) tma_store_wait_read<0>();
      asm volatile("bar.sync %0, 128;" ::"r"(1 + quarter) : "memory");
      if (warp == 0 && lane == 0) stamp(it, 2);
      if (kResidual && active && elected) {
        mbar_expect_tx(rfull_bar(quarter), OUT_BYTES);
        tma_load_2d(buf, &tmap_r, rfull_bar(quarter), colq, row_base);
      }
      float2 st = make_float2(0.f, 1.f);  // (mean, rstd) of this row
      if ((kLnIn || kLnRes) && row_ok) {
        if (ln.part_in) {
          // hidden is a multiple of 128, so a row holds an even number (<= 16) of 8-byte slots: fetch
          // them as float4 pairs, all loads in flight together, then add in slot order (every reader
          // gets the same bits)
          const float4* pp = reinterpret_cast<const float4*>(ln.part_in + static_cast<size_t>(row) * ln.part_slots);
          const int pairs = ln.part_slots >> 1;
          float4 pv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) pv[i] = i < pairs ? __ldg(pp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
          float s = 0.f, sq = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (i < pairs) {
              s += pv[i].x; sq += pv[i].y;
              s += pv[i].z; sq += pv[i].w;
            }
          }
          const float mean = s * ln.inv_h;
          const float var = fmaxf(sq * ln.inv_h - mean * mean, 0.f);
          st = make_float2(mean, rsqrtf(var + ln.eps));
        } else {
          st = __ldg(ln.row_stats + row);
        }
      }
      float psum = 0.f, psq = 0.f;
      mbar_wait(tfull_bar(as), aphase);
      if (warp == 0 && lane == 0) stamp(it, 3);
      tc_fence_after();
      if (kResidual && active) {
        mbar_wait(rfull_bar(quarter), rphase);
        rphase ^= 1u;
      }
      if (warp == 0 && lane == 0) stamp(it, 4);
      const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + quarter * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tbase + c * 32, r);
        tmem_ld_wait();
        if (c == 1) {
          // both halves are in registers: hand the accumulator stage back before the arithmetic
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(tempty_bar(as), 0);  // the issuer waits on the leader's copy
          if (warp == 0 && lane == 0) stamp(it, 5);
        }
        if (active) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int col = colq + c * 32 + j * 8;
            uint4* slot = reinterpret_cast<uint4*>(buf_ptr + (((c * 4 + j) ^ sw) << 4));
            // columns at or beyond N are clipped by the TMA store; keep their loads in bounds
            const int colc = col < N ? col : 0;
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + colc));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + colc + 4));
            float v[8];
            if (kLnIn) {
              // acc·r − μ·r·g + c   (the bias slot carries c)
              const float4 g0 = __ldg(reinterpret_cast<const float4*>(ln.g + colc));
              const float4 g1 = __ldg(reinterpret_cast<const float4*>(ln.g + colc + 4));
              // packed fp32x2: two columns per instruction, same operations and rounding as scalar fmaf
              const float mr = st.x * st.y;
              const uint64_t nmr2 = f32x2(-mr, -mr), rs2 = f32x2(st.y, st.y);
              const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
              const float cc[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
              for (int e = 0; e < 8; e += 2) {
                const uint64_t t = ffma2(nmr2, f32x2(gg[e], gg[e + 1]), f32x2(cc[e], cc[e + 1]));
                f32x2_split(ffma2(f32x2_bits(r[j * 8 + e], r[j * 8 + e + 1]), rs2, t), v[e], v[e + 1]);
              }
            } else {
              v[0] = __uint_as_float(r[j * 8 + 0]) + b0.x;
              v[1] = __uint_as_float(r[j * 8 + 1]) + b0.y;
              v[2] = __uint_as_float(r[j * 8 + 2]) + b0.z;
              v[3] = __uint_as_float(r[j * 8 + 3]) + b0.w;
              v[4] = __uint_as_float(r[j * 8 + 4]) + b1.x;
              v[5] = __uint_as_float(r[j * 8 + 5]) + b1.y;
              v[6] = __uint_as_float(r[j * 8 + 6]) + b1.z;
              v[7] = __uint_as_float(r[j * 8 + 7]) + b1.w;
            }
            if (EPI == GEMM_EPI_BIAS_RESIDUAL) {
              const uint4 rv = *slot;
              v[0] += bf16_lo(rv.x); v[1] += bf16_hi(rv.x);
              v[2] += bf16_lo(rv.y); v[3] += bf16_hi(rv.y);
              v[4] += bf16_lo(rv.z); v[5] += bf16_hi(rv.z);
              v[6] += bf16_lo(rv.w); v[7] += bf16_hi(rv.w);
            }
            if (kLnRes) {
              // residual = LayerNorm(U) rebuilt from the un-normalised row: (U − μ)·r·γ + β
              const uint4 rv = *slot;
              const float4 ga0 = __ldg(reinterpret_cast<const float4*>(ln.res_gamma + colc));
              const float4 ga1 = __ldg(reinterpret_cast<const float4*>(ln.res_gamma + colc + 4));
              const float4 be0 = __ldg(reinterpret_cast<const float4*>(ln.res_beta + colc));
              const float4 be1 = __ldg(reinterpret_cast<const float4*>(ln.res_beta + colc + 4));
              v[0] += fmaf((bf16_lo(rv.x) - st.x) * st.y, ga0.x, be0.x);
              v[1] += fmaf((bf16_hi(rv.x) - st.x) * st.y, ga0.y, be0.y);
              v[2] += fmaf((bf16_lo(rv.y) - st.x) * st.y, ga0.z, be0.z);
              v[3] += fmaf((bf16_hi(rv.y) - st.x) * st.y, ga0.w, be0.w);
              v[4] += fmaf((bf16_lo(rv.z) - st.x) * st.y, ga1.x, be1.x);
              v[5] += fmaf((bf16_hi(rv.z) - st.x) * st.y, ga1.y, be1.y);
              v[6] += fmaf((bf16_lo(rv.w) - st.x) * st.y, ga1.z, be1.z);
              v[7] += fmaf((bf16_hi(rv.w) - st.x) * st.y, ga1.w, be1.w);
              if (col < N) {
#pragma unroll
                for (int e = 0; e < 8; ++e) { psum += v[e]; psq = fmaf(v[e], v[e], psq); }
              }
            }
            if (EPI == GEMM_EPI_BIAS_GELU || EPI == GEMM_EPI_LNIN_GELU) {
#pragma unroll
              for (int e = 0; e < 8; e += 2) gelu_erf2(v[e], v[e + 1]);
            } else {
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = epi_act<EPI>(v[e]);
            }
            // Rows at or beyond M inside the last row tile are stored too (the TMA store clips at the
            // buffer, not at M). They must stay finite: they are derived from whatever the same rows
            // held before, batch after batch and layer after layer, and once such a feedback loop
            // reaches Inf the NaN that follows poisons real rows through masked-out attention products
            // (0 x NaN) — it took ~100 batches of BERT-base to get there. Zeros end the loop.
            uint4 o = make_uint4(0u, 0u, 0u, 0u);
            if (row_ok) {
              o.x = pack_bf16x2(v[0], v[1]);
              o.y = pack_bf16x2(v[2], v[3]);
              o.z = pack_bf16x2(v[4], v[5]);
              o.w = pack_bf16x2(v[6], v[7]);
            }
            *slot = o;
          }
        }
      }
      if (kLnRes && row_ok && active)
        ln.part_out[static_cast<size_t>(row) * ln.part_slots + (colq >> 6)] = make_float2(psum, psq);
      if (warp == 0 && lane == 0) stamp(it, 6);
      fence_proxy_async_smem();  // staging block written through the generic proxy, read by TMA
      asm volatile("bar.sync %0, 128;" ::"r"(1 + quarter) : "memory");
      if (active && elected) {
        tma_store_2d(&tmap_c, buf, colq, row_base);  // rows >= rows-of-C and columns >= N are clipped
        tma_store_commit();
      }
      if (warp == 0 && lane == 0) stamp(it, 7);
    }
    if (elected) tma_store_wait_all<0>();  // the staging block must outlive its last store
  }

  __syncwarp();  // reconverge the single-thread roles before the aligned cluster barrier
  tc_fence_before();
  cluster_sync_all();  // neither CTA may free TMEM (or exit) while the other still uses the pair
  if (warp == W_ALLOC) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, 512);
  }
}

unsigned long long* g_gemm2_trace = nullptr;
int g_gemm2_debug = 0;

template <int EPI>
int launch2(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
            const CUtensorMap& tr, const float* bias, int M, int N, int K, int num_sms,
            const int32_t* m_dev, const GemmLnArgs& ln) {
  static bool configured[64] = {};
  int dev = 0;
  B2_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2_CUDA_OK(cudaFuncSetAttribute(gemm2_bf16_tn_kernel<EPI>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    configured[dev] = true;
  }
  const int tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + BN - 1) / BN);
  int pairs = num_sms / 2;
  if (tiles < pairs) pairs = tiles;
  GemmLnArgs args = ln;
  args.trace = g_gemm2_trace;
  args.debug_mode = g_gemm2_debug;
  gemm2_bf16_tn_kernel<EPI><<<2 * pairs, THREADS, SMEM_BYTES, stream>>>(ta, tb, tc, tr, bias, M, N, K,
                                                                       m_dev, args);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace

void gemm2_set_trace(unsigned long long* dev_buf, int debug_mode) {
  g_gemm2_trace = dev_buf;
  g_gemm2_debug = debug_mode;
}

// W tensor map for the pair kernel: each CTA loads 128 rows of W per K step.
int gemm2_tmap_b(CUtensorMap* out, const void* W, int N, int K) {
  return make_tmap_bf16_2d(out, W, N, K, static_cast<uint64_t>(K) * 2, BNH, BK);
}

// Output / residual tensor map: [rows, cols] bf16 row-major, 128 x 64 boxes, 128-byte swizzle (the
// same geometry as an A-operand map, so one map per activation buffer serves both purposes).
int gemm2_tmap_c(CUtensorMap* out, const void* Cbuf, int rows, int cols, int64_t ldc) {
  return make_tmap_bf16_2d(out, Cbuf, rows, cols, static_cast<uint64_t>(ldc) * 2, 128, 64);
}

int gemm2_bf16_tn(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                  const CUtensorMap* tr, const float* bias, int M, int N, int K, int epilogue, int num_sms,
                  const int32_t* m_dev, const GemmLnArgs* ln_in) {
  const GemmLnArgs ln = ln_in ? *ln_in : GemmLnArgs();
  if (M <= 0 || N <= 0 || K <= 0 || (N % 8) != 0 || (K % 8) != 0 || num_sms < 2 || !bias) {
    set_last_error("gemm2_bf16_tn: bad shape M=%d N=%d K=%d (N, K must be multiples of 8; bias required)", M,
                   N, K);
    return B2_ERR_ARG;
  }
  const bool needs_res = epilogue == GEMM_EPI_BIAS_RESIDUAL || epilogue == GEMM_EPI_LNRES_STATS;
  if (needs_res && !tr) {
    set_last_error("gemm2_bf16_tn: residual epilogue needs the residual's tensor map");
    return B2_ERR_ARG;
  }
  const CUtensorMap& trr = tr ? *tr : tc;
  switch (epilogue) {
    case GEMM_EPI_BIAS:
      return launch2<GEMM_EPI_BIAS>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
    case GEMM_EPI_BIAS_GELU:
      return launch2<GEMM_EPI_BIAS_GELU>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
    case GEMM_EPI_BIAS_RESIDUAL:
      return launch2<GEMM_EPI_BIAS_RESIDUAL>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
    case GEMM_EPI_BIAS_TANH:
      return launch2<GEMM_EPI_BIAS_TANH>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
    case GEMM_EPI_LNIN:
    case GEMM_EPI_LNIN_GELU:
      if (ln.part_in && (ln.part_slots < 2 || ln.part_slots > 16 || (ln.part_slots & 1))) {
        set_last_error("gemm2_bf16_tn: part_in needs an even number of slots in [2, 16]");
        return B2_ERR_ARG;
      }
      if ((!ln.row_stats && !ln.part_in) || !ln.g) {
        set_last_error("gemm2_bf16_tn: LNIN epilogue needs row_stats and g");
        return B2_ERR_ARG;
      }
      return epilogue == GEMM_EPI_LNIN
                 ? launch2<GEMM_EPI_LNIN>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln)
                 : launch2<GEMM_EPI_LNIN_GELU>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
    case GEMM_EPI_LNRES_STATS:
      if ((!ln.row_stats && !ln.part_in) || !ln.res_gamma || !ln.res_beta || !ln.part_out || ln.part_slots * 64 < N) {
        set_last_error("gemm2_bf16_tn: LNRES_STATS epilogue needs row_stats, gamma, beta and part_out with "
                       "part_slots >= N/64");
        return B2_ERR_ARG;
      }
      return launch2<GEMM_EPI_LNRES_STATS>(stream, ta, tb, tc, trr, bias, M, N, K, num_sms, m_dev, ln);
  }
  set_last_error("gemm2_bf16_tn: unknown epilogue %d", epilogue);
  return B2_ERR_ARG;
}

int gemm2_set_trap_log(unsigned long long* mapped) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_trap_log, &mapped, sizeof(mapped)));
  return B2_OK;
}

}  // namespace b2
