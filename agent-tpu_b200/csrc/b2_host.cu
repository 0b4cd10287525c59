// b2_host.cu — host-side helpers: thread-local error string, TMA descriptor encoding.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "b2_kernels.h"

namespace b2 {

static thread_local char g_last_error[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_last_error; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn resolve_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                      uint64_t pitch_bytes, uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn encode = resolve_encode();
  if (!encode) {
    set_last_error("cuTensorMapEncodeTiled is not available from this driver");
    return B2_ERR_CUDA;
  }
  if (box_cols * 2 != 128 || box_rows > 256 || (pitch_bytes % 16) != 0 ||
      (reinterpret_cast<uintptr_t>(base) % 16) != 0) {
    set_last_error("make_tmap_bf16_2d: unsupported box %ux%u / pitch %llu / base alignment", box_rows,
                   box_cols, (unsigned long long)pitch_bytes);
    return B2_ERR_ARG;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t elem_strides[2] = {1, 1};
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                      strides, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu pitch=%llu)",
                   (int)r, (unsigned long long)rows, (unsigned long long)cols,
                   (unsigned long long)pitch_bytes);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

}  // namespace b2
