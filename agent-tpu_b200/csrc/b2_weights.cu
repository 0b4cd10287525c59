// b2_weights.cu — weight materialisation on the device: the deterministic synthetic initialiser
// (bit-identical to model_file.synthetic_tensor on the host) and f32 -> bf16 conversion.
//
// Reference counterpart: model loading in get_tpu_handle (ops/_tpu_runtime.py:55-56,
// make_interpreter + allocate_tensors). The reference's model file is absent from its repo
// (ops/_tpu_runtime.py:27-31), so the "file" here may describe its tensors by seed instead.
#include "b2_kernels.h"

namespace b2 {

namespace {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}

// value(i) = scale * (l0 + l1 + l2 + l3 - 131070) + shift, l* the four 16-bit lanes of the hash.
// Integer sum, one f32 multiply, one f32 add: no transcendental, so numpy reproduces it exactly.
__global__ void fill_synth_kernel(float* __restrict__ dst, int64_t n, uint64_t seed,
                                  uint32_t tensor_id, float scale, float shift) {
  const uint64_t key = seed ^ (static_cast<uint64_t>(tensor_id) * 0xD1B54A32D192ED03ull);
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const uint64_t z = splitmix64(key + (static_cast<uint64_t>(i) + 1ull) * 0x9E3779B97F4A7C15ull);
    const int s = static_cast<int>((z & 0xFFFF) + ((z >> 16) & 0xFFFF) + ((z >> 32) & 0xFFFF) +
                                   (z >> 48)) - 131070;
    dst[i] = __fadd_rn(__fmul_rn(scale, static_cast<float>(s)), shift);
  }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                   int64_t n) {
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// one warp per output row n of W[N,K]
__global__ void fold_ln_kernel(const float* __restrict__ W, const float* __restrict__ bias,
                               const float* __restrict__ gamma, const float* __restrict__ beta, int N, int K,
                               __nv_bfloat16* __restrict__ Wf, float* __restrict__ g, float* __restrict__ c) {
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (n >= N) return;
  const int lane = threadIdx.x & 31;
  const float* w = W + static_cast<size_t>(n) * K;
  __nv_bfloat16* wf = Wf + static_cast<size_t>(n) * K;
  float gs = 0.f, cs = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float wk = w[k];
    const __nv_bfloat16 f = __float2bfloat16_rn(gamma ? gamma[k] * wk : wk);
    wf[k] = f;
    gs += __bfloat162float(f);  // the sum of what the tensor core will actually multiply
    if (beta) cs = fmaf(beta[k], wk, cs);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gs += __shfl_xor_sync(0xffffffffu, gs, o);
    cs += __shfl_xor_sync(0xffffffffu, cs, o);
  }
  if (lane == 0) {
    g[n] = gs;
    c[n] = bias[n] + cs;
  }
}

}  // namespace

int fold_ln_into_linear(cudaStream_t stream, const float* W, const float* bias, const float* gamma,
                        const float* beta, int N, int K, __nv_bfloat16* Wf, float* g, float* c) {
  if (N <= 0 || K <= 0) return B2_OK;
  fold_ln_kernel<<<(N + 7) / 8, 256, 0, stream>>>(W, bias, gamma, beta, N, K, Wf, g, c);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int fill_synthetic_f32(cudaStream_t stream, float* dst, int64_t n, uint64_t seed,
                       uint32_t tensor_id, float scale, float shift) {
  if (n <= 0) return B2_OK;
  const int blocks = static_cast<int>(n / 256 + 1 < 148 * 16 ? n / 256 + 1 : 148 * 16);
  fill_synth_kernel<<<blocks, 256, 0, stream>>>(dst, n, seed, tensor_id, scale, shift);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int convert_f32_to_bf16(cudaStream_t stream, const float* src, __nv_bfloat16* dst, int64_t n) {
  if (n <= 0) return B2_OK;
  const int blocks = static_cast<int>(n / 256 + 1 < 148 * 16 ? n / 256 + 1 : 148 * 16);
  f32_to_bf16_kernel<<<blocks, 256, 0, stream>>>(src, dst, n);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace b2
