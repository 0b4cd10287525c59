// b2_gemm.cu — persistent warp-specialised bf16 GEMM on tcgen05 tensor cores, operands by TMA.
//
//   C[M,N] (bf16) = epilogue( A[M,K] (bf16, row-major) · W[N,K]^T (bf16, row-major) + bias[N] )
//
// This is the contraction the reference delegates to `interpreter.invoke()`
// (ops/map_classify_tpu.py:71-72); here it carries the encoder's QKV, attention-output, FFN1 and
// FFN2 projections and the pooler (SURVEY.md §8(a′)).
//
// Tile plan (one CTA per SM, persistent over tiles):
//   CTA tile 128 x 256, K step 64 (= one 128-byte swizzle atom of bf16), 4-stage TMA ring
//   (48 KiB / stage), one tcgen05.mma of 128x256x16 per K=16, f32 accumulators in TMEM,
//   two accumulator stages (2 x 256 columns = all 512 TMEM columns) so the epilogue of tile i
//   overlaps the main loop of tile i+1.
// Warp roles (640 threads): warp 16 = TMA producer, warp 17 = MMA issuer, warp 18 = TMEM allocator,
//   warp 19 idle, warps 0..15 = epilogue (TMEM lane quadrant = warp % 4, 64-column quarter =
//   warp/4). Sixteen epilogue warps (four per SM sub-partition) because the epilogue is a chain
//   of dependent FP32/MUFU operations per element: with two warps per sub-partition it ran at 0.35
//   IPC and paced the FFN1 GEMM (profiles/r01_ncu_gemm.md).
#include "b2_kernels.h"

namespace b2 {

namespace {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int STAGES = 4;
constexpr int THREADS = 640;
constexpr int EPI_WARPS = 16;
// Warp roles. The SM sub-partition arbiter prefers the highest warp id among eligible warps, so the
// single-thread roles that pace the tensor pipe (TMA producer, MMA issuer) sit ABOVE the sixteen
// epilogue warps: placed below them they were starved of issue slots whenever the epilogue was busy,
// and every instruction added to the epilogue slowed the main loop down.
constexpr int W_TMA = 16, W_MMA = 17, W_ALLOC = 18;
constexpr uint32_t A_BYTES = BM * BK * 2;  // 16 KiB
constexpr uint32_t B_BYTES = BN * BK * 2;  // 32 KiB
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t BAR_BYTES = 8 * (2 * STAGES + 4) + 16;
constexpr uint32_t SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + BAR_BYTES;

template <int EPI>
__device__ __forceinline__ float epi_act(float v) {
  if (EPI == GEMM_EPI_BIAS_GELU) return gelu_erf(v);
  if (EPI == GEMM_EPI_BIAS_TANH) return tanhf(v);
  return v;
}

template <int EPI>
__global__ void __launch_bounds__(THREADS, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b, __nv_bfloat16* __restrict__ C,
                    const float* __restrict__ bias, const __nv_bfloat16* __restrict__ residual,
                    int M, int N, int K, int ldc, int ldr, const int32_t* __restrict__ m_dev) {
  extern __shared__ uint8_t smem_raw[];
  // packed (padding-free) batches: the row count is produced on the device by the pack kernel, so
  // the host never has to synchronise to learn it; `M` is then only the upper bound
  if (m_dev) M = min(M, __ldg(m_dev));
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t sA = base;
  const uint32_t sB = base + STAGES * A_BYTES;
  const uint32_t bars = base + STAGES * STAGE_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (M + BM - 1) / BM;
  const int n_tiles = (N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == W_TMA && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp == W_MMA && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), EPI_WARPS);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == W_ALLOC) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == W_TMA) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int m_blk = t / n_tiles, n_blk = t % n_tiles;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), STAGE_BYTES);
          tma_load_2d(sA + stage * A_BYTES, &tmap_a, full_bar(stage), kb * BK, m_blk * BM);
          tma_load_2d(sB + stage * B_BYTES, &tmap_b, full_bar(stage), kb * BK, n_blk * BN);
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == W_MMA) {
    // ===== MMA issuer (single thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_sw128(sA + stage * A_BYTES, 16, 1024);
          const uint64_t b_desc = umma_desc_sw128(sB + stage * B_BYTES, 16, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // +32 bytes per K=16 step inside the 128-byte swizzle atom: +2 in the >>4 address field
            umma_bf16(d_tmem, a_desc + 2u * k, b_desc + 2u * k, idesc, (kb | k) != 0);
          }
          umma_commit(empty_bar(stage));  // frees the smem slot once these MMAs have read it
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(as));  // accumulator complete -> epilogue
      }
    }
  } else if (warp < EPI_WARPS) {
    // ===== epilogue: TMEM -> registers -> bias / activation / residual -> bf16 -> global =====
    const int q = warp & 3;               // TMEM lane quadrant this warp may read
    const int quarter = warp >> 2;  // 64-column quarter of the 256-wide tile
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
      const int m_blk = t / n_tiles, n_blk = t % n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const int row = m_blk * BM + q * 32 + lane;
      const bool row_ok = row < M;
      const int colq = n_blk * BN + quarter * 64;
      __nv_bfloat16* crow = C + static_cast<size_t>(row) * ldc;
      // the residual row segment does not depend on the accumulator: fetch it before waiting
      uint4 res[8];
      if (EPI == GEMM_EPI_BIAS_RESIDUAL) {
        const __nv_bfloat16* rrow = residual + static_cast<size_t>(row) * ldr;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int col = colq + j * 8;
          res[j] = (row_ok && col < N) ? __ldg(reinterpret_cast<const uint4*>(rrow + col))
                                       : make_uint4(0u, 0u, 0u, 0u);
        }
      }
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + quarter * 64;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tbase + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int col = colq + c * 32 + j * 8;
          if (row_ok && col < N) {
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + col));
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + col + 4));
            float v[8];
            v[0] = __uint_as_float(r[j * 8 + 0]) + b0.x;
            v[1] = __uint_as_float(r[j * 8 + 1]) + b0.y;
            v[2] = __uint_as_float(r[j * 8 + 2]) + b0.z;
            v[3] = __uint_as_float(r[j * 8 + 3]) + b0.w;
            v[4] = __uint_as_float(r[j * 8 + 4]) + b1.x;
            v[5] = __uint_as_float(r[j * 8 + 5]) + b1.y;
            v[6] = __uint_as_float(r[j * 8 + 6]) + b1.z;
            v[7] = __uint_as_float(r[j * 8 + 7]) + b1.w;
            if (EPI == GEMM_EPI_BIAS_RESIDUAL) {
              const uint4 rv = res[c * 4 + j];
              v[0] += bf16_lo(rv.x); v[1] += bf16_hi(rv.x);
              v[2] += bf16_lo(rv.y); v[3] += bf16_hi(rv.y);
              v[4] += bf16_lo(rv.z); v[5] += bf16_hi(rv.z);
              v[6] += bf16_lo(rv.w); v[7] += bf16_hi(rv.w);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = epi_act<EPI>(v[e]);
            uint4 o;
            o.x = pack_bf16x2(v[0], v[1]);
            o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]);
            o.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(crow + col) = o;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == W_ALLOC) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

template <int EPI>
int launch(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb, __nv_bfloat16* C,
           const float* bias, const __nv_bfloat16* residual, int M, int N, int K, int ldc, int ldr,
           int num_sms, const int32_t* m_dev) {
  static bool configured[64] = {};
  int dev = 0;
  B2_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    B2_CUDA_OK(cudaFuncSetAttribute(gemm_bf16_tn_kernel<EPI>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    configured[dev] = true;
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < num_sms ? tiles : num_sms;
  gemm_bf16_tn_kernel<EPI><<<grid, THREADS, SMEM_BYTES, stream>>>(ta, tb, C, bias, residual, M, N,
                                                                  K, ldc, ldr, m_dev);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace

int gemm_tmap_a(CUtensorMap* out, const void* A, int M, int K, int64_t lda) {
  return make_tmap_bf16_2d(out, A, M, K, static_cast<uint64_t>(lda) * 2, BM, BK);
}
int gemm_tmap_b(CUtensorMap* out, const void* W, int N, int K) {
  return make_tmap_bf16_2d(out, W, N, K, static_cast<uint64_t>(K) * 2, BN, BK);
}

int gemm_bf16_tn(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb,
                 __nv_bfloat16* C, const float* bias, const __nv_bfloat16* residual, int M, int N,
                 int K, int ldc, int ldr, int epilogue, int num_sms, const int32_t* m_dev) {
  if (M <= 0 || N <= 0 || K <= 0 || (N % 8) != 0 || (K % 8) != 0 || (ldc % 8) != 0) {
    set_last_error("gemm_bf16_tn: bad shape M=%d N=%d K=%d ldc=%d (N, K, ldc must be multiples of 8)",
                   M, N, K, ldc);
    return B2_ERR_ARG;
  }
  switch (epilogue) {
    case GEMM_EPI_BIAS:
      return launch<GEMM_EPI_BIAS>(stream, ta, tb, C, bias, residual, M, N, K, ldc, ldr, num_sms, m_dev);
    case GEMM_EPI_BIAS_GELU:
      return launch<GEMM_EPI_BIAS_GELU>(stream, ta, tb, C, bias, residual, M, N, K, ldc, ldr,
                                        num_sms, m_dev);
    case GEMM_EPI_BIAS_RESIDUAL:
      if (!residual || (ldr % 8) != 0) {
        set_last_error("gemm_bf16_tn: residual epilogue needs a residual pointer and ldr %% 8 == 0");
        return B2_ERR_ARG;
      }
      return launch<GEMM_EPI_BIAS_RESIDUAL>(stream, ta, tb, C, bias, residual, M, N, K, ldc, ldr,
                                            num_sms, m_dev);
    case GEMM_EPI_BIAS_TANH:
      return launch<GEMM_EPI_BIAS_TANH>(stream, ta, tb, C, bias, residual, M, N, K, ldc, ldr,
                                        num_sms, m_dev);
  }
  set_last_error("gemm_bf16_tn: unknown epilogue %d", epilogue);
  return B2_ERR_ARG;
}

int gemm_set_trap_log(unsigned long long* mapped) {
  B2_CUDA_OK(cudaMemcpyToSymbol(g_trap_log, &mapped, sizeof(mapped)));
  return B2_OK;
}

}  // namespace b2
