// b2_pace.cu — developer micro-benchmark (dev_check tmapace), not on any product path: how fast can the
// SMs pull the pair GEMM's operand stream out of L2, and does TMA multicast change that?
//
// The pair GEMM's main loop needs 32 KiB per SM per 64-deep K step (its CTA's 128 rows of A and 128 rows
// of W). With the MMAs switched off that stream alone runs at ~565 cycles per step (dev_check
// gemm2trace), above the 512 cycles the tensor pipe needs — 58 B/clk/SM, 12.9 TB/s for the chip, which is
// what the L2 can deliver. This kernel reproduces the stream with the same tensor maps, tile order and
// five-stage ring and nothing else, in two forms:
//   mode 0  every CTA loads its own A tile (16 KiB) and W half (16 KiB): what the GEMM does today;
//   mode 1  clusters of four CTAs = two CTA pairs working on the same column tile: a CTA loads its A
//           tile and one QUARTER of the W tile (8 KiB) and multicasts it to the CTA of the other pair
//           that needs the same half, so 24 KiB per CTA and step are requested from L2 instead of 32.
// All tensor maps cover buffers of this GPU only (TMA multicast loads here read only this GPU's memory,
// since on B200 a TMA multicast load from another GPU's memory hung the GPU).
#include "b2_kernels.h"

namespace b2 {
namespace {

constexpr uint32_t P_TILE = 128 * 64 * 2;  // 16 KiB
constexpr uint32_t p_smem(int stages) { return 1024 + stages * 2 * P_TILE + 256; }

__device__ __forceinline__ uint32_t p_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t p_clusterid() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void p_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void p_arrive_remote(uint32_t bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, %1;\n\tmbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
      ::"r"(bar), "r"(rank) : "memory");
}
// the same box lands at the same shared-memory offset of every CTA in `mask`, and each of them gets the
// bytes counted on its own barrier at the same offset
__device__ __forceinline__ void p_tma_load_mc(uint32_t dst, const CUtensorMap* tmap, uint32_t bar, int32_t c0,
                                              int32_t c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
}

// hold > 0: the consumer keeps every slot for `hold` cycles, in order, the way the tensor pipe keeps a
// stage until the MMAs that read it have run (512 cycles per 64-deep step at peak): the ring then has
// to cover the TMA latency PLUS that time, which is the situation of the real main loop.
template <int MODE, int P_STAGES>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(128, 1)
tma_pace_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                const __grid_constant__ CUtensorMap tmap_wq, int m_tiles, int n_tiles, int num_kb, int tiles,
                int hold, unsigned long long* __restrict__ cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + P_STAGES * 2 * P_TILE;
  auto full = [&](int s) { return bars + 8u * s; };
  auto empty = [&](int s) { return bars + 8u * (P_STAGES + s); };
  const uint32_t rank = p_ctarank(), pair = rank >> 1, half = rank & 1, peer = rank ^ 2u;
  const int pairs_total = (gridDim.x / 4) * 2;
  const int my_pair = p_clusterid() * 2 + pair;
  if (threadIdx.x == 0) {
    for (int s = 0; s < P_STAGES; ++s) {
      mbar_init(full(s), 1);
      mbar_init(empty(s), MODE == 1 ? 2 : 1);
    }
    fence_barrier_init();
  }
  p_cluster_sync();
  if (threadIdx.x == 0) {
    // producer
    const long long t0 = clock64();
    int stage = 0;
    uint32_t phase = 0;
    for (int t = 0; t < tiles; ++t) {
      const int m_blk = (my_pair + t * pairs_total) % m_tiles;
      const int n_blk = (p_clusterid() + t) % n_tiles;   // the cluster's two pairs share the column tile
      const int arow = m_blk * 256 + half * 128;
      const int wrow = n_blk * 256 + half * 128;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(empty(stage), phase ^ 1u);
        mbar_expect_tx(full(stage), 2 * P_TILE);
        const uint32_t sA = base + stage * 2 * P_TILE, sW = sA + P_TILE;
        tma_load_2d(sA, &tmap_a, full(stage), kb * 64, arow);
        if (MODE == 0) {
          tma_load_2d(sW, &tmap_w, full(stage), kb * 64, wrow);
        } else {
          p_tma_load_mc(sW + pair * (P_TILE / 2), &tmap_wq, full(stage), kb * 64, wrow + pair * 64,
                        static_cast<uint16_t>((1u << rank) | (1u << peer)));
        }
        if (++stage == P_STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    cycles[blockIdx.x * 2] = static_cast<unsigned long long>(clock64() - t0);
  } else if (threadIdx.x == 32) {
    // consumer: stands in for the MMA thread — the slot is free as soon as its bytes have landed
    const long long t0 = clock64();
    int stage = 0;
    uint32_t phase = 0;
    long long busy_until = t0;
    for (int i = 0; i < tiles * num_kb; ++i) {
      mbar_wait(full(stage), phase);
      if (hold > 0) {
        const long long now = clock64();
        busy_until = (now > busy_until ? now : busy_until) + hold;
        while (clock64() < busy_until) {}
      }
      mbar_arrive(empty(stage));
      if (MODE == 1) p_arrive_remote(empty(stage), peer);  // the peer multicasts into this slot too
      if (++stage == P_STAGES) { stage = 0; phase ^= 1u; }
    }
    cycles[blockIdx.x * 2 + 1] = static_cast<unsigned long long>(clock64() - t0);
  }
  __syncwarp();
  p_cluster_sync();  // no CTA may leave while a peer can still write into it
}

}  // namespace

// A: [rows, K] bf16, W: [n, K] bf16, both in this device's memory. cycles: device buffer of 2 * grid u64.
template <int MODE, int STAGES>
static int launch_pace(cudaStream_t stream, int grid, const CUtensorMap& ta, const CUtensorMap& tw,
                       const CUtensorMap& twq, int m_tiles, int n_tiles, int num_kb, int tiles, int hold,
                       unsigned long long* cycles) {
  B2_CUDA_OK(cudaFuncSetAttribute(tma_pace_kernel<MODE, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  p_smem(STAGES)));
  tma_pace_kernel<MODE, STAGES><<<grid, 128, p_smem(STAGES), stream>>>(ta, tw, twq, m_tiles, n_tiles, num_kb, tiles,
                                                                       hold, cycles);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int tma_pace(cudaStream_t stream, int mode, int stages, int hold, const void* A, int rows, const void* W, int n,
             int K, int tiles, int num_sms, unsigned long long* cycles, int* grid_out) {
  CUtensorMap ta, tw, twq;
  int rc;
  if ((rc = make_tmap_bf16_2d(&ta, A, rows, K, static_cast<uint64_t>(K) * 2, 128, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&tw, W, n, K, static_cast<uint64_t>(K) * 2, 128, 64))) return rc;
  if ((rc = make_tmap_bf16_2d(&twq, W, n, K, static_cast<uint64_t>(K) * 2, 64, 64))) return rc;
  // 32 clusters of four: all resident at once on 148 SMs whatever the GPCs' sizes, same grid for both modes
  const int grid = num_sms / 4 * 4 < 128 ? num_sms / 4 * 4 : 128;
  const int m_tiles = rows / 256, n_tiles = n / 256, num_kb = K / 64;
  if (grid_out) *grid_out = grid;
  if (mode == 1) return launch_pace<1, 5>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
  switch (stages) {
    case 3: return launch_pace<0, 3>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
    case 4: return launch_pace<0, 4>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
    case 5: return launch_pace<0, 5>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
    case 6: return launch_pace<0, 6>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
    case 7: return launch_pace<0, 7>(stream, grid, ta, tw, twq, m_tiles, n_tiles, num_kb, tiles, hold, cycles);
  }
  set_last_error("tma_pace: stages must be 3..7");
  return B2_ERR_ARG;
}

}  // namespace b2
