// dev_check.cu — developer self-check binary (not part of the shipped library): runs each kernel
// on random data and compares with a straightforward host computation. Built by `make dev_check`.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "b2_kernels.h"

using namespace b2;

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e = (x);                                                               \
    if (e != cudaSuccess) {                                                            \
      printf("CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e)); \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)
#define B2(x)                                                        \
  do {                                                               \
    int r = (x);                                                     \
    if (r != 0) {                                                    \
      printf("b2 error %d at %s:%d: %s\n", r, __FILE__, __LINE__, last_error()); \
      exit(3);                                                       \
    }                                                                \
  } while (0)

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static float frand() {
  rng_state = rng_state * 6364136223846793005ull + 1442695040888963407ull;
  return ((rng_state >> 40) & 0xFFFFFF) / float(1 << 24) * 2.0f - 1.0f;
}
static float bf16_round(float f) { return __bfloat162float(__float2bfloat16(f)); }

static double gelu_ref(double x) { return 0.5 * x * (1.0 + erf(x / sqrt(2.0))); }

static int check_gemm(int M, int N, int K, int epi, int num_sms, bool time_it) {
  std::vector<__nv_bfloat16> hA((size_t)M * K), hW((size_t)N * K), hR((size_t)M * N);
  std::vector<float> fA((size_t)M * K), fW((size_t)N * K), fR((size_t)M * N), hb(N);
  for (size_t i = 0; i < hA.size(); ++i) { fA[i] = bf16_round(frand()); hA[i] = __float2bfloat16(fA[i]); }
  for (size_t i = 0; i < hW.size(); ++i) { fW[i] = bf16_round(frand() * 0.05f); hW[i] = __float2bfloat16(fW[i]); }
  for (size_t i = 0; i < hR.size(); ++i) { fR[i] = bf16_round(frand()); hR[i] = __float2bfloat16(fR[i]); }
  for (int i = 0; i < N; ++i) hb[i] = frand() * 0.5f;
  __nv_bfloat16 *dA, *dW, *dC, *dR;
  float* db;
  CK(cudaMalloc(&dA, hA.size() * 2));
  CK(cudaMalloc(&dW, hW.size() * 2));
  CK(cudaMalloc(&dC, (size_t)M * N * 2));
  CK(cudaMalloc(&dR, (size_t)M * N * 2));
  CK(cudaMalloc(&db, N * 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dW, hW.data(), hW.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dR, hR.data(), hR.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), N * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dC, 0xFF, (size_t)M * N * 2));
  CUtensorMap ta, tb;
  B2(gemm_tmap_a(&ta, dA, M, K, K));
  B2(gemm_tmap_b(&tb, dW, N, K));
  B2(gemm_bf16_tn(0, ta, tb, dC, db, dR, M, N, K, N, N, epi, num_sms));
  CK(cudaDeviceSynchronize());
  std::vector<__nv_bfloat16> hC((size_t)M * N);
  CK(cudaMemcpy(hC.data(), dC, hC.size() * 2, cudaMemcpyDeviceToHost));
  // check: all entries for small problems, a strided sample otherwise
  double max_err = 0;
  size_t bad = 0, checked = 0;
  size_t total = (size_t)M * N;
  size_t step = total > (1u << 20) ? total / (1u << 18) + 1 : 1;
  for (size_t idx = 0; idx < total; idx += step) {
    int m = idx / N, n = idx % N;
    double acc = 0;
    for (int k = 0; k < K; ++k) acc += (double)fA[(size_t)m * K + k] * fW[(size_t)n * K + k];
    acc += hb[n];
    if (epi == GEMM_EPI_BIAS_RESIDUAL) acc += fR[idx];
    if (epi == GEMM_EPI_BIAS_GELU) acc = gelu_ref(acc);
    if (epi == GEMM_EPI_BIAS_TANH) acc = tanh(acc);
    double got = __bfloat162float(hC[idx]);
    double err = fabs(got - acc);
    double tol = 0.01 * fabs(acc) + 0.01;
    if (!(err <= tol)) { if (bad < 5) printf("  mismatch m=%d n=%d got=%f want=%f\n", m, n, got, acc); ++bad; }
    if (err > max_err) max_err = err;
    ++checked;
  }
  printf("gemm M=%d N=%d K=%d epi=%d: checked=%zu bad=%zu max_abs_err=%.5f %s\n", M, N, K, epi,
         checked, bad, max_err, bad ? "FAIL" : "ok");
  if (time_it && !bad) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) B2(gemm_bf16_tn(0, ta, tb, dC, db, dR, M, N, K, N, N, epi, num_sms));
    CK(cudaEventRecord(e0));
    const int iters = 10;
    for (int i = 0; i < iters; ++i) B2(gemm_bf16_tn(0, ta, tb, dC, db, dR, M, N, K, N, N, epi, num_sms));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= iters;
    printf("  time %.3f ms  %.1f TFLOP/s\n", ms, 2.0 * M * N * K / ms / 1e9);
  }
  cudaFree(dA); cudaFree(dW); cudaFree(dC); cudaFree(dR); cudaFree(db);
  return bad ? 1 : 0;
}

int main(int argc, char** argv) {
  const char* what = argc > 1 ? argv[1] : "all";
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  printf("device: %s sm_%d%d, %d SMs\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  int sms = prop.multiProcessorCount;
  int fails = 0;
  if (!strcmp(what, "all") || !strcmp(what, "gemm")) {
    fails += check_gemm(128, 256, 64, GEMM_EPI_BIAS, sms, false);
    fails += check_gemm(128, 256, 768, GEMM_EPI_BIAS, sms, false);
    fails += check_gemm(256, 512, 128, GEMM_EPI_BIAS, sms, false);
    fails += check_gemm(200, 264, 72, GEMM_EPI_BIAS, sms, false);      // ragged M, N, K
    fails += check_gemm(1000, 768, 3072, GEMM_EPI_BIAS_RESIDUAL, sms, false);
    fails += check_gemm(1024, 3072, 768, GEMM_EPI_BIAS_GELU, sms, false);
    fails += check_gemm(512, 768, 768, GEMM_EPI_BIAS_TANH, sms, false);
    fails += check_gemm(40000, 2304, 768, GEMM_EPI_BIAS, sms, false);   // > 148 tiles: persistence
  }
  if (!strcmp(what, "all") || !strcmp(what, "gemmperf")) {
    fails += check_gemm(131072, 2304, 768, GEMM_EPI_BIAS, sms, true);
    fails += check_gemm(131072, 768, 768, GEMM_EPI_BIAS_RESIDUAL, sms, true);
    fails += check_gemm(131072, 3072, 768, GEMM_EPI_BIAS_GELU, sms, true);
    fails += check_gemm(131072, 768, 3072, GEMM_EPI_BIAS_RESIDUAL, sms, true);
  }
  printf("%s (%d failing checks)\n", fails ? "DEV_CHECK FAILED" : "DEV_CHECK PASSED", fails);
  return fails ? 1 : 0;
}
