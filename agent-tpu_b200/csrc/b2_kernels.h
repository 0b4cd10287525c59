// b2_kernels.h — internal launcher interface between the C-ABI layer (b2_api.cu) and the kernels.
// Nothing here is exported; the exported surface is include/b2_classify.h.
#pragma once
#include "b2_common.cuh"

namespace b2 {

// ---- GEMM (b2_gemm.cu) --------------------------------------------------------------------------
enum GemmEpilogue {
  GEMM_EPI_BIAS = 0,           // C = A·Wᵀ + b
  GEMM_EPI_BIAS_GELU = 1,      // C = gelu_erf(A·Wᵀ + b)
  GEMM_EPI_BIAS_RESIDUAL = 2,  // C = A·Wᵀ + b + R
  GEMM_EPI_BIAS_TANH = 3,      // C = tanh(A·Wᵀ + b)      (pooler)
};
// Tensor maps: A is [M,K] bf16 with row pitch lda elements, W is [N,K] bf16 dense.
int gemm_tmap_a(CUtensorMap* out, const void* A, int M, int K, int64_t lda);
int gemm_tmap_b(CUtensorMap* out, const void* W, int N, int K);
// m_dev (optional, device memory): actual row count, M is then an upper bound (packed batches).
int gemm_bf16_tn(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb,
                 __nv_bfloat16* C, const float* bias, const __nv_bfloat16* residual, int M, int N,
                 int K, int ldc, int ldr, int epilogue, int num_sms, const int32_t* m_dev = nullptr);

// Same contract on CTA pairs (b2_gemm2.cu: cta_group::2, 256 x 256 pair tile). `ta` is the same A map;
// `tb` must come from gemm2_tmap_b (128-row boxes of W).
int gemm2_tmap_b(CUtensorMap* out, const void* W, int N, int K);
int gemm2_bf16_tn(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb,
                  __nv_bfloat16* C, const float* bias, const __nv_bfloat16* residual, int M, int N,
                  int K, int ldc, int ldr, int epilogue, int num_sms, const int32_t* m_dev = nullptr);

// ---- attention (b2_attention.cu) ---------------------------------------------------------------
// qkv: [rows, 3*H] bf16 (Q | K | V per token row, head h at columns h*64), out: [rows, H] bf16.
// lens[b] = number of valid (unpadded) tokens of sequence b; keys >= lens[b] are masked out and
// query rows >= lens[b] are not written. Sequence b starts at row cu[b] (packed) or b*S (cu null).
// Four tensor maps over the same buffer (TMA boxes of 32, 64, 96 and 128 token rows): the kernel
// picks the smallest box that covers a sequence's valid rows.
struct AttentionMaps {
  alignas(64) unsigned char raw[4 * sizeof(CUtensorMap)];
};
int attention_tmaps(AttentionMaps* out, const void* qkv, int rows, int hidden);
int attention_bf16(cudaStream_t stream, const AttentionMaps& tqkv, __nv_bfloat16* out,
                   const int32_t* lens, int B, int S, int hidden, const int32_t* cu = nullptr);

// ---- element-wise / row kernels (b2_rows.cu) ---------------------------------------------------
// out[row] = LayerNorm(word[ids[row]] + pos[row % S] + type[0]) * gamma + beta  -> bf16
// packed: with lens / cu given, token (b, s) is written to row cu[b] + s and pad positions are skipped.
int embed_layernorm(cudaStream_t stream, const int32_t* ids, const float* word, const float* pos,
                    const float* type0, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int rows, int S, int hidden, int vocab,
                    const int32_t* lens = nullptr, const int32_t* cu = nullptr);
// x[row] = LayerNorm(x[row]) * gamma + beta   (in place, bf16 storage, f32 math)
int layernorm_inplace(cudaStream_t stream, __nv_bfloat16* x, const float* gamma, const float* beta,
                      float eps, int rows, int hidden, const int32_t* rows_dev = nullptr);
// cu[0..B] = exclusive prefix sum of lens, cu[B] = total rows (also the GEMMs' m_dev); B <= 4096.
int pack_offsets(cudaStream_t stream, const int32_t* lens, int B, int32_t* cu);
// dst[b, :] = src[first row of sequence b, :]  (the [CLS] rows): row cu[b] (packed) or b*S (cu null)
int gather_cls_rows(cudaStream_t stream, const __nv_bfloat16* src, const int32_t* cu, int S, int B,
                    int hidden, __nv_bfloat16* dst);
// logits[b, c] = pooled[b,:] · Wc[c,:] + bc[c]; probs = softmax(logits); top-k (descending, ties
// by lower index), label = argmax, histogram[label] += 1.
int classify_head(cudaStream_t stream, const __nv_bfloat16* pooled, int ld_pooled, const float* Wc,
                  const float* bc, int B, int hidden, int C, int topk, float* logits,
                  int32_t* top_idx, float* top_score, int32_t* labels,
                  unsigned long long* histogram);
// token counts per sequence from the mask: lens[b] = sum(mask[b,:])
int mask_to_lens(cudaStream_t stream, const uint8_t* mask, int32_t* lens, int B, int S);

// ---- weights (b2_weights.cu) -------------------------------------------------------------------
// Deterministic synthetic initialiser shared bit-for-bit with the host generator (model_file.py):
// value(i) = scale * (sum of four 16-bit lanes of splitmix64(seed, tensor_id, i) - 131070) + shift
int fill_synthetic_f32(cudaStream_t stream, float* dst, int64_t n, uint64_t seed,
                       uint32_t tensor_id, float scale, float shift);
int convert_f32_to_bf16(cudaStream_t stream, const float* src, __nv_bfloat16* dst, int64_t n);

}  // namespace b2
