// b2_rows.cu — the HBM-bound row kernels around the GEMMs: embedding gather + LayerNorm,
// LayerNorm, the classifier head (logits, softmax, top-k, argmax, label histogram).
//
// Reference counterparts: the head replaces `_topk` (ops/map_classify_tpu.py:15-19: clamp k to
// [1,n], descending by score) and the float32 read-back at ops/map_classify_tpu.py:74; embedding
// and LayerNorm have none (BertEmbeddings / BertSelfOutput / BertOutput semantics, eps inside sqrt).
//
// One warp per row everywhere: 16- or 8-byte vector loads, f32 math, shuffle reductions.
#include "b2_kernels.h"

namespace b2 {

namespace {

constexpr int MAX_VEC = 8;  // hidden <= 8 * 128

// ---- embedding gather + LayerNorm -------------------------------------------------------------
__global__ void __launch_bounds__(256)
embed_ln_kernel(const int32_t* __restrict__ ids, const float* __restrict__ word,
                const float* __restrict__ pos, const float* __restrict__ type0,
                const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                __nv_bfloat16* __restrict__ out, int rows, int S, int hidden, int vocab,
                const int32_t* __restrict__ lens, const int32_t* __restrict__ cu) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);  // padded position b*S + s
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  int out_row = row;
  if (cu) {
    const int b = row / S, sp = row - b * S;
    if (sp >= lens[b]) return;  // pad position: no row in the packed layout
    out_row = cu[b] + sp;
  }
  int id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float* w = word + static_cast<size_t>(id) * hidden;
  const float* p = pos + static_cast<size_t>(row % S) * hidden;
  const int nvec = hidden >> 7;
  float4 v[MAX_VEC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const int c = i * 128 + lane * 4;
      const float4 a = __ldg(reinterpret_cast<const float4*>(w + c));
      const float4 b = __ldg(reinterpret_cast<const float4*>(p + c));
      const float4 t = __ldg(reinterpret_cast<const float4*>(type0 + c));
      v[i] = make_float4(a.x + b.x + t.x, a.y + b.y + t.y, a.z + b.z + t.z, a.w + b.w + t.w);
      sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(sum) / hidden;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const float dx = v[i].x - mean, dy = v[i].y - mean, dz = v[i].z - mean, dw = v[i].w - mean;
      sq += (dx * dx + dy * dy) + (dz * dz + dw * dw);
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / hidden + eps);
  __nv_bfloat16* o = out + static_cast<size_t>(out_row) * hidden;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const int c = i * 128 + lane * 4;
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c));
      const float4 bb = __ldg(reinterpret_cast<const float4*>(beta + c));
      uint2 w2;
      w2.x = pack_bf16x2((v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y);
      w2.y = pack_bf16x2((v[i].z - mean) * rstd * g.z + bb.z, (v[i].w - mean) * rstd * g.w + bb.w);
      *reinterpret_cast<uint2*>(o + c) = w2;
    }
  }
}

// ---- LayerNorm in place over bf16 rows ----------------------------------------------------------
__global__ void __launch_bounds__(256)
layernorm_kernel(__nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                 const float* __restrict__ beta, float eps, int rows, int hidden,
                 const int32_t* __restrict__ rows_dev) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (rows_dev) rows = min(rows, __ldg(rows_dev));
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  __nv_bfloat16* xr = x + static_cast<size_t>(row) * hidden;
  const int nvec = hidden >> 7;
  float4 v[MAX_VEC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const uint2 raw = *reinterpret_cast<const uint2*>(xr + i * 128 + lane * 4);
      v[i] = make_float4(bf16_lo(raw.x), bf16_hi(raw.x), bf16_lo(raw.y), bf16_hi(raw.y));
      sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(sum) / hidden;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const float dx = v[i].x - mean, dy = v[i].y - mean, dz = v[i].z - mean, dw = v[i].w - mean;
      sq += (dx * dx + dy * dy) + (dz * dz + dw * dw);
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / hidden + eps);
#pragma unroll
  for (int i = 0; i < MAX_VEC; ++i) {
    if (i < nvec) {
      const int c = i * 128 + lane * 4;
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c));
      const float4 bb = __ldg(reinterpret_cast<const float4*>(beta + c));
      uint2 w2;
      w2.x = pack_bf16x2((v[i].x - mean) * rstd * g.x + bb.x, (v[i].y - mean) * rstd * g.y + bb.y);
      w2.y = pack_bf16x2((v[i].z - mean) * rstd * g.z + bb.z, (v[i].w - mean) * rstd * g.w + bb.w);
      *reinterpret_cast<uint2*>(xr + c) = w2;
    }
  }
}

// ---- classifier head ----------------------------------------------------------------------------
// One warp per item. logits in shared memory (C floats per warp), then softmax, top-k by repeated
// arg-max (k and C are small: the contract's default is k = 5), label = arg-max, histogram.
constexpr int HEAD_WARPS = 4;

__global__ void __launch_bounds__(HEAD_WARPS * 32)
classify_head_kernel(const __nv_bfloat16* __restrict__ pooled, int ld_pooled,
                     const float* __restrict__ Wc, const float* __restrict__ bc, int B, int hidden,
                     int C, int topk, float* __restrict__ logits_out, int32_t* __restrict__ top_idx,
                     float* __restrict__ top_score, int32_t* __restrict__ labels,
                     unsigned long long* __restrict__ histogram) {
  extern __shared__ float head_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int item = blockIdx.x * HEAD_WARPS + warp;
  if (item >= B) return;
  float* lg = head_smem + warp * C;
  const __nv_bfloat16* x = pooled + static_cast<size_t>(item) * ld_pooled;

  for (int c = 0; c < C; ++c) {
    const float* w = Wc + static_cast<size_t>(c) * hidden;
    float acc = 0.f;
    for (int k = lane * 4; k < hidden; k += 128) {
      const uint2 raw = *reinterpret_cast<const uint2*>(x + k);
      const float4 ww = __ldg(reinterpret_cast<const float4*>(w + k));
      acc = fmaf(bf16_lo(raw.x), ww.x, acc);
      acc = fmaf(bf16_hi(raw.x), ww.y, acc);
      acc = fmaf(bf16_lo(raw.y), ww.z, acc);
      acc = fmaf(bf16_hi(raw.y), ww.w, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) lg[c] = acc + bc[c];
  }
  __syncwarp();
  if (logits_out)
    for (int c = lane; c < C; c += 32) logits_out[static_cast<size_t>(item) * C + c] = lg[c];

  // softmax statistics
  float mx = -INFINITY;
  for (int c = lane; c < C; c += 32) mx = fmaxf(mx, lg[c]);
  mx = warp_max(mx);
  float den = 0.f;
  for (int c = lane; c < C; c += 32) den += __expf(lg[c] - mx);
  den = warp_sum(den);
  const float inv = 1.0f / den;

  // top-k: k rounds of (max value, lowest index) over the not-yet-taken classes
  for (int t = 0; t < topk; ++t) {
    float best = -INFINITY;
    int best_i = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = lg[c];
      if (v > best || (v == best && c < best_i)) { best = v; best_i = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
      if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
    }
    if (lane == 0) {
      top_idx[static_cast<size_t>(item) * topk + t] = best_i;
      top_score[static_cast<size_t>(item) * topk + t] = __expf(best - mx) * inv;
      if (t == 0) {
        labels[item] = best_i;
        if (histogram) atomicAdd(histogram + best_i, 1ull);
      }
      lg[best_i] = -INFINITY;  // logits were already copied out above
    }
    __syncwarp();
  }
}

// exclusive prefix sum of lens[0..B) -> cu[0..B], one CTA of 1024 threads, up to 4 items per thread
__global__ void __launch_bounds__(1024) pack_offsets_kernel(const int32_t* __restrict__ lens, int B,
                                                            int32_t* __restrict__ cu) {
  __shared__ int warp_tot[32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  int v[4], sum = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = t * 4 + i;
    v[i] = idx < B ? lens[idx] : 0;
    sum += v[i];
  }
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = warp_tot[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += y;
    }
    warp_tot[lane] = w;  // inclusive totals of warps 0..lane
  }
  __syncthreads();
  int run = incl - sum + (warp > 0 ? warp_tot[warp - 1] : 0);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = t * 4 + i;
    if (idx < B) cu[idx] = run;
    run += v[i];
  }
  if (t == 1023) cu[B] = warp_tot[31];
}

__global__ void gather_cls_kernel(const __nv_bfloat16* __restrict__ src, const int32_t* __restrict__ cu,
                                  int S, int B, int hidden, __nv_bfloat16* __restrict__ dst) {
  const int b = blockIdx.x;
  if (b >= B) return;
  const size_t row = cu ? static_cast<size_t>(cu[b]) : static_cast<size_t>(b) * S;
  const uint2* s = reinterpret_cast<const uint2*>(src + row * hidden);
  uint2* d = reinterpret_cast<uint2*>(dst + static_cast<size_t>(b) * hidden);
  for (int i = threadIdx.x; i < hidden / 4; i += blockDim.x) d[i] = s[i];
}

// ---- LayerNorm statistics for the folded path ---------------------------------------------------
__global__ void ln_stats_finalize_kernel(const float2* __restrict__ part, int slots, int rows, float inv_h,
                                         float eps, float2* __restrict__ stats,
                                         const int32_t* __restrict__ rows_dev) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (rows_dev) rows = min(rows, __ldg(rows_dev));
  if (row >= rows) return;
  const float2* p = part + static_cast<size_t>(row) * slots;
  float s = 0.f, q = 0.f;
  for (int i = 0; i < slots; ++i) {  // fixed order: the sums are reproducible
    const float2 v = p[i];
    s += v.x;
    q += v.y;
  }
  const float mean = s * inv_h;
  const float var = fmaxf(q * inv_h - mean * mean, 0.f);
  stats[row] = make_float2(mean, rsqrtf(var + eps));
}

__global__ void ln_stats_identity_kernel(float2* __restrict__ stats, int rows) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row < rows) stats[row] = make_float2(0.f, 1.f);
}

__global__ void gather_cls_norm_kernel(const __nv_bfloat16* __restrict__ src, const float2* __restrict__ stats,
                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                       const int32_t* __restrict__ cu, int S, int B, int hidden,
                                       __nv_bfloat16* __restrict__ dst) {
  const int b = blockIdx.x;
  if (b >= B) return;
  const size_t row = cu ? static_cast<size_t>(cu[b]) : static_cast<size_t>(b) * S;
  const float2 st = stats[row];
  const __nv_bfloat16* s = src + row * hidden;
  __nv_bfloat16* d = dst + static_cast<size_t>(b) * hidden;
  for (int c = threadIdx.x * 2; c < hidden; c += blockDim.x * 2) {
    const uint32_t raw = *reinterpret_cast<const uint32_t*>(s + c);
    const float a = fmaf((bf16_lo(raw) - st.x) * st.y, gamma[c], beta[c]);
    const float e = fmaf((bf16_hi(raw) - st.x) * st.y, gamma[c + 1], beta[c + 1]);
    *reinterpret_cast<uint32_t*>(d + c) = pack_bf16x2(a, e);
  }
}

// lens[b] = number of ones in mask row b. Every later kernel treats a sequence as the prefix
// [0, lens[b]) of its row, so a row is only acceptable if its ones ARE that prefix and there is at
// least one: the lowest offending row number is left in *bad_row (atomicMin; caller presets it high).
__global__ void mask_to_lens_kernel(const uint8_t* __restrict__ mask, int32_t* __restrict__ lens,
                                    int B, int S, int32_t* __restrict__ bad_row) {
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  const int lane = threadIdx.x & 31;
  int n = 0, last = -1;
  for (int s = lane; s < S; s += 32) {
    if (mask[static_cast<size_t>(b) * S + s] != 0) { ++n; last = s; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n += __shfl_xor_sync(0xffffffffu, n, o);
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  }
  if (lane == 0) {
    lens[b] = n;
    if (bad_row && (n < 1 || last + 1 != n)) atomicMin(bad_row, b);
  }
}

// ---- developer aid: hidden states as the consumers see them ----------------------------------------
// out[(b*S + s)*H + c] = float of the layer input the next kernel would read for token (b, s):
// (x − μ)·r·γ + β with (μ, r) from finished statistics, from the partial sums (added in slot order,
// like every production reader), or the identity when x is already normalised. One warp per token.
__global__ void __launch_bounds__(256)
dump_hidden_kernel(const __nv_bfloat16* __restrict__ x, const float2* __restrict__ stats,
                   const float2* __restrict__ part, int slots, float inv_h, float eps,
                   const float* __restrict__ gamma, const float* __restrict__ beta,
                   const int32_t* __restrict__ lens, const int32_t* __restrict__ cu, int B, int S, int H,
                   float* __restrict__ out) {
  const int pos = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pos >= B * S) return;
  const int lane = threadIdx.x & 31;
  const int b = pos / S, sp = pos - b * S;
  if (sp >= lens[b]) return;
  const size_t row = cu ? static_cast<size_t>(cu[b]) + sp : static_cast<size_t>(pos);
  float mean = 0.f, rstd = 1.f;
  if (part) {
    const float2* p = part + row * slots;
    float s = 0.f, q = 0.f;
    for (int i = 0; i < slots; ++i) { const float2 v = p[i]; s += v.x; q += v.y; }
    mean = s * inv_h;
    rstd = rsqrtf(fmaxf(q * inv_h - mean * mean, 0.f) + eps);
  } else if (stats) {
    const float2 st = stats[row];
    mean = st.x; rstd = st.y;
  }
  const __nv_bfloat16* xr = x + row * H;
  float* o = out + static_cast<size_t>(pos) * H;
  for (int c = lane; c < H; c += 32) {
    const float v = (__bfloat162float(xr[c]) - mean) * rstd;
    o[c] = gamma ? fmaf(v, gamma[c], beta[c]) : v;
  }
}

// out[(b*S)*H + c] = float(cls[b*H + c]): the last layer only exists for the [CLS] rows
__global__ void dump_cls_kernel(const __nv_bfloat16* __restrict__ cls, int B, int S, int H,
                                float* __restrict__ out) {
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < H; c += blockDim.x)
    out[static_cast<size_t>(b) * S * H + c] = __bfloat162float(cls[static_cast<size_t>(b) * H + c]);
}

}  // namespace

int embed_layernorm(cudaStream_t stream, const int32_t* ids, const float* word, const float* pos,
                    const float* type0, const float* gamma, const float* beta, float eps,
                    __nv_bfloat16* out, int rows, int S, int hidden, int vocab, const int32_t* lens,
                    const int32_t* cu) {
  if ((hidden % 128) != 0 || hidden > MAX_VEC * 128) {
    set_last_error("embed_layernorm: hidden must be a multiple of 128 and <= %d", MAX_VEC * 128);
    return B2_ERR_ARG;
  }
  if (rows <= 0) return B2_OK;
  const int wpb = 8;
  embed_ln_kernel<<<(rows + wpb - 1) / wpb, wpb * 32, 0, stream>>>(ids, word, pos, type0, gamma, beta,
                                                                  eps, out, rows, S, hidden, vocab, lens, cu);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int layernorm_inplace(cudaStream_t stream, __nv_bfloat16* x, const float* gamma, const float* beta,
                      float eps, int rows, int hidden, const int32_t* rows_dev) {
  if ((hidden % 128) != 0 || hidden > MAX_VEC * 128) {
    set_last_error("layernorm_inplace: hidden must be a multiple of 128 and <= %d", MAX_VEC * 128);
    return B2_ERR_ARG;
  }
  if (rows <= 0) return B2_OK;
  const int wpb = 8;
  layernorm_kernel<<<(rows + wpb - 1) / wpb, wpb * 32, 0, stream>>>(x, gamma, beta, eps, rows, hidden,
                                                                   rows_dev);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int classify_head(cudaStream_t stream, const __nv_bfloat16* pooled, int ld_pooled, const float* Wc,
                  const float* bc, int B, int hidden, int C, int topk, float* logits,
                  int32_t* top_idx, float* top_score, int32_t* labels,
                  unsigned long long* histogram) {
  if (C <= 0 || C > 2048 || topk < 1 || topk > C || (hidden % 4) != 0 || (ld_pooled % 4) != 0) {
    set_last_error("classify_head: need 1 <= topk <= C <= 2048 and hidden %% 4 == 0 (C=%d topk=%d)", C,
                   topk);
    return B2_ERR_ARG;
  }
  if (B <= 0) return B2_OK;
  const size_t smem = static_cast<size_t>(HEAD_WARPS) * C * sizeof(float);
  classify_head_kernel<<<(B + HEAD_WARPS - 1) / HEAD_WARPS, HEAD_WARPS * 32, smem, stream>>>(
      pooled, ld_pooled, Wc, bc, B, hidden, C, topk, logits, top_idx, top_score, labels, histogram);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int pack_offsets(cudaStream_t stream, const int32_t* lens, int B, int32_t* cu) {
  if (B < 0 || B > 4096) {
    set_last_error("pack_offsets: batch of %d sequences exceeds 4096", B);
    return B2_ERR_LIMIT;
  }
  pack_offsets_kernel<<<1, 1024, 0, stream>>>(lens, B, cu);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int gather_cls_rows(cudaStream_t stream, const __nv_bfloat16* src, const int32_t* cu, int S, int B,
                    int hidden, __nv_bfloat16* dst) {
  if (B <= 0) return B2_OK;
  gather_cls_kernel<<<B, 128, 0, stream>>>(src, cu, S, B, hidden, dst);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int ln_stats_finalize(cudaStream_t stream, const float2* part, int slots, int rows, int hidden, float eps,
                      float2* stats, const int32_t* rows_dev) {
  if (rows <= 0) return B2_OK;
  ln_stats_finalize_kernel<<<(rows + 255) / 256, 256, 0, stream>>>(part, slots, rows, 1.0f / hidden, eps,
                                                                    stats, rows_dev);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int ln_stats_identity(cudaStream_t stream, float2* stats, int rows) {
  if (rows <= 0) return B2_OK;
  ln_stats_identity_kernel<<<(rows + 255) / 256, 256, 0, stream>>>(stats, rows);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int gather_cls_rows_norm(cudaStream_t stream, const __nv_bfloat16* src, const float2* stats,
                         const float* gamma, const float* beta, const int32_t* cu, int S, int B, int hidden,
                         __nv_bfloat16* dst) {
  if (B <= 0) return B2_OK;
  gather_cls_norm_kernel<<<B, 128, 0, stream>>>(src, stats, gamma, beta, cu, S, B, hidden, dst);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int mask_to_lens(cudaStream_t stream, const uint8_t* mask, int32_t* lens, int B, int S, int32_t* bad_row) {
  if (B <= 0) return B2_OK;
  mask_to_lens_kernel<<<(B + 7) / 8, 256, 0, stream>>>(mask, lens, B, S, bad_row);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int dump_hidden(cudaStream_t stream, const __nv_bfloat16* x, const float2* stats, const float2* part, int slots,
                float eps, const float* gamma, const float* beta, const int32_t* lens, const int32_t* cu,
                int B, int S, int H, float* out) {
  if (B <= 0) return B2_OK;
  const int rows = B * S;
  dump_hidden_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(x, stats, part, slots, 1.0f / H, eps, gamma, beta, lens,
                                                         cu, B, S, H, out);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

int dump_cls(cudaStream_t stream, const __nv_bfloat16* cls, int B, int S, int H, float* out) {
  if (B <= 0) return B2_OK;
  dump_cls_kernel<<<B, 256, 0, stream>>>(cls, B, S, H, out);
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}

}  // namespace b2
