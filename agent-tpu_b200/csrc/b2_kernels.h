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
  // LayerNorm folded into the GEMMs around it (pair kernel only, see GemmLnArgs):
  GEMM_EPI_LNIN = 4,           // C = r[m]·(A·W'ᵀ − μ[m]·g) + c            A holds un-normalised rows
  GEMM_EPI_LNIN_GELU = 5,      // C = gelu_erf(r[m]·(A·W'ᵀ − μ[m]·g) + c)
  GEMM_EPI_LNRES_STATS = 6,    // C = A·Wᵀ + b + ((U − μ)·r·γ + β); per-row Σ, Σ² of C -> part_out
};
// Arguments of the LayerNorm-folding epilogues. With x = (U − μ)·r·γ + β never materialised:
//   x·Wᵀ + b = r·(U·(γ∘W)ᵀ − μ·g) + c,   g[n] = Σ_k (γ∘W)[n,k],   c[n] = b[n] + Σ_k β[k]·W[n,k]
// so a GEMM can consume the un-normalised rows U directly (LNIN), and the residual add of the
// GEMM that closes a sub-layer can rebuild x from U on the fly while it emits the row sums the next
// LayerNorm needs (LNRES_STATS). Row statistics travel as float2 (mean, rstd) per row.
struct GemmLnArgs {
  const float2* row_stats = nullptr;  // LNIN: stats of A's rows. LNRES: stats of the residual rows U
  // ... or, instead of finished statistics, the partial sums an LNRES_STATS epilogue left behind
  // ([rows, part_slots] (Σ, Σ²)); every reader adds them in slot order, so all readers agree
  const float2* part_in = nullptr;
  float inv_h = 0.f, eps = 0.f;       // 1 / row length and LayerNorm epsilon (with part_in)
  const float* g = nullptr;           // LNIN: column sums of the folded weight (bias slot carries c)
  const float* res_gamma = nullptr;   // LNRES: γ, β of the LayerNorm that produces the residual
  const float* res_beta = nullptr;
  float2* part_out = nullptr;         // LNRES: [rows, part_slots] partial (Σ, Σ²), slot = column/64
  int part_slots = 0;                 // slots per row of part_in and part_out
  unsigned long long* trace = nullptr;  // developer timeline (dev_check gemm2trace), else null
  int debug_mode = 0;  // dev_check only: 1 = no TMA loads after the first ring fill, 2 = no MMAs (results garbage)
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
// `tc` (output) and `tr` (residual, may be null) are 128 x 64-box maps of [rows, cols] bf16 buffers
// (gemm2_tmap_c, same geometry as gemm_tmap_a): the epilogue stores through TMA and fetches the
// residual through TMA, so rows beyond the buffer and columns beyond N are clipped by hardware.
int gemm2_tmap_c(CUtensorMap* out, const void* Cbuf, int rows, int cols, int64_t ldc);
int gemm2_bf16_tn(cudaStream_t stream, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                  const CUtensorMap* tr, const float* bias, int M, int N, int K, int epilogue, int num_sms,
                  const int32_t* m_dev = nullptr, const GemmLnArgs* ln = nullptr);

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

void attention_set_trace(unsigned long long* dev_buf);  // developer timeline hook (dev_check)
void gemm2_set_trace(unsigned long long* dev_buf, int debug_mode = 0);  // same for the pair GEMM

// developer micro-benchmark (b2_pace.cu, dev_check tmapace): the pair GEMM's operand stream alone,
// unicast (mode 0) or with the W tile multicast across a cluster of two CTA pairs (mode 1)
// `stages` = ring depth (3..7, mode 0); hold > 0: the consumer keeps each slot that many cycles, in order
int tma_pace(cudaStream_t stream, int mode, int stages, int hold, const void* A, int rows, const void* W, int n,
             int K, int tiles, int num_sms, unsigned long long* cycles, int* grid_out);

// developer aid: bounded waits that give up record where (b2_common.cuh: mbar_wait_timeout)
int gemm_set_trap_log(unsigned long long* mapped);
int gemm2_set_trap_log(unsigned long long* mapped);
int attention_set_trap_log(unsigned long long* mapped);

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
// stats[row] = (mean, rsqrt(var + eps)) from the partial sums a GEMM_EPI_LNRES_STATS epilogue wrote;
// partials are added in slot order, so the result does not depend on scheduling.
int ln_stats_finalize(cudaStream_t stream, const float2* part, int slots, int rows, int hidden, float eps,
                      float2* stats, const int32_t* rows_dev = nullptr);
// stats[row] = (0, 1): rows that are already normalised (the embedding output)
int ln_stats_identity(cudaStream_t stream, float2* stats, int rows);
// dst[b,:] = (src[first row of sequence b,:] − μ)·r·γ + β   (gather the [CLS] rows and normalise them)
int gather_cls_rows_norm(cudaStream_t stream, const __nv_bfloat16* src, const float2* stats,
                         const float* gamma, const float* beta, const int32_t* cu, int S, int B, int hidden,
                         __nv_bfloat16* dst);
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
// token counts per sequence from the mask: lens[b] = sum(mask[b,:]). The encoder treats sequence b as
// the prefix [0, lens[b]) of its row; *bad_row (device, preset to INT_MAX by the caller, may be null)
// receives the lowest row whose ones are not exactly such a non-empty prefix.
int mask_to_lens(cudaStream_t stream, const uint8_t* mask, int32_t* lens, int B, int S,
                 int32_t* bad_row = nullptr);
// developer aid (b2_debug_hidden_states): f32 copy of a layer's input rows as the consuming kernels
// see them — (x − μ)·r·γ + β with statistics from `part` (partial sums) or `stats`, identity if both
// are null — into out[B, S, H] at padded positions; pad positions are left untouched.
int dump_hidden(cudaStream_t stream, const __nv_bfloat16* x, const float2* stats, const float2* part, int slots,
                float eps, const float* gamma, const float* beta, const int32_t* lens, const int32_t* cu,
                int B, int S, int H, float* out);
int dump_cls(cudaStream_t stream, const __nv_bfloat16* cls, int B, int S, int H, float* out);

// ---- CSV (b2_csv.cu) ------------------------------------------------------------------------------
// Record splitter with the semantics of Python's csv reader (default dialect) over bytes
// [first, n) of a buffer; the data starts a record at `first`. `scratch` needs
// csv_index_scratch_bytes(n) bytes and must stay untouched between csv_count_rows and
// csv_emit_rows. `bytes` must be readable up to the next multiple of 16.
size_t csv_index_scratch_bytes(int64_t n_bytes);
// d_total_rows: number of non-empty records; d_final_state: automaton state after the last byte
// (anything but 0 means the last row has no line end and ends at n); *d_first_bad = first offset in
// [0, n) at which strict UTF-8 decoding fails (n: data ends inside a sequence), 2^64-1 if valid.
int csv_count_rows(cudaStream_t st, const uint8_t* bytes, int64_t first, int64_t n, void* scratch,
                   int64_t* d_total_rows, int32_t* d_final_state, unsigned long long* d_first_bad);
// row_start[r] = offset of the first byte of row r, row_end[r] = offset of the line end closing it
// (not written for a last row without line end).
int csv_emit_rows(cudaStream_t st, int64_t first, int64_t n, void* scratch, int64_t* row_start,
                  int64_t* row_end);
// d_out2[0] = first row holding a field of more than `limit` code points, d_out2[1] = offset of
// the code point that overflows it; both 2^64-1 when there is none.
int csv_field_limit(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                    int64_t n_rows, int64_t limit, unsigned long long* d_out2);
// Rows [r0, r0 + n_rows): number of fields and bytes of unquoted field content per row. col >= 0
// keeps only that column (one field per row, empty when the row is shorter).
int csv_rows_size(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                  int64_t r0, int64_t n_rows, int32_t col, int64_t* n_fields, int64_t* n_bytes);
// Same walk, writing: content of row i at out_bytes + byte_base[i], end offset (in out_bytes) of its
// field f at field_end[field_base[i] + f] (col < 0) or of its one field at field_end[i] (col >= 0).
int csv_rows_write(cudaStream_t st, const uint8_t* bytes, const int64_t* row_start, const int64_t* row_end,
                   int64_t r0, int64_t n_rows, int32_t col, const int64_t* field_base,
                   const int64_t* byte_base, uint8_t* out_bytes, int64_t* field_end);
// out[0..n] = exclusive prefix sum of in[0..n), out[n] = total. scratch: scan_i64_scratch_bytes(n).
size_t scan_i64_scratch_bytes(int64_t n);
int exclusive_scan_i64(cudaStream_t st, const int64_t* in, int64_t n, int64_t* out, void* scratch);

// ---- weights (b2_weights.cu) -------------------------------------------------------------------
// Deterministic synthetic initialiser shared bit-for-bit with the host generator (model_file.py):
// value(i) = scale * (sum of four 16-bit lanes of splitmix64(seed, tensor_id, i) - 131070) + shift
int fill_synthetic_f32(cudaStream_t stream, float* dst, int64_t n, uint64_t seed,
                       uint32_t tensor_id, float scale, float shift);
int convert_f32_to_bf16(cudaStream_t stream, const float* src, __nv_bfloat16* dst, int64_t n);
// Fold a LayerNorm's affine into the Linear that follows it: Wf[n,k] = bf16(gamma[k]·W[n,k]),
// g[n] = Σ_k float(Wf[n,k]), c[n] = bias[n] + Σ_k beta[k]·W[n,k]. gamma/beta null = identity.
int fold_ln_into_linear(cudaStream_t stream, const float* W, const float* bias, const float* gamma,
                        const float* beta, int N, int K, __nv_bfloat16* Wf, float* g, float* c);

}  // namespace b2
