// b2_api.cu — the exported C ABI (include/b2_classify.h): handle life cycle, model file loading,
// and the per-job orchestration of the kernels (tokenize -> embed+LN -> 12 x {QKV GEMM, attention,
// out-proj GEMM + residual, LN, FFN1 GEMM + GELU, FFN2 GEMM + residual, LN} -> pooler GEMM + tanh
// -> head).
//
// It is the B200 counterpart of get_tpu_handle (ops/_tpu_runtime.py:34-63: load once, cache, check
// the file exists) and of the set_tensor / invoke / get_tensor / _topk sequence in
// ops/map_classify_tpu.py:61-75, widened from one tensor per call to a whole job of items.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "b2_tokenizer.h"

namespace b2 {

namespace {

// ---- model file (.b2m) ---------------------------------------------------------------------------
// header, little endian (written by agent-tpu_b200/model_file.py):
//   char magic[4] = "B2MD"; u32 version = 1;
//   u32 hidden, layers, heads, intermediate, vocab_size, max_positions, type_vocab, num_labels;
//   f32 ln_eps; u32 kind (0 = f32 tensors follow, 1 = synthetic); u64 seed;
//   f32 init_std, bias_std, gamma_std, classifier_std;
// kind 0: the tensors follow in the order of `tensor_list()` below, f32.
// kind 1: tensor t, element i = scale_t * (sum of 16-bit lanes of splitmix64(seed, t, i) - 131070)
//         + shift_t, with scale_t = std_t / sqrt(4 * (65536^2 - 1) / 12)  (see b2_weights.cu).
struct ModelHeader {
  char magic[4];
  uint32_t version;
  uint32_t hidden, layers, heads, intermediate, vocab_size, max_positions, type_vocab, num_labels;
  float ln_eps;
  uint32_t kind;
  uint64_t seed;
  float init_std, bias_std, gamma_std, classifier_std;
};
static_assert(sizeof(ModelHeader) == 72, "model header layout");

enum TensorRole { ROLE_WEIGHT, ROLE_BIAS, ROLE_GAMMA, ROLE_BETA, ROLE_CLASSIFIER };

struct TensorSpec {
  uint32_t id;     // stable tensor id (hash stream selector for synthetic models)
  int64_t n;       // elements
  TensorRole role;
};

// ids: 0 word, 1 pos, 2 type, 3 emb_ln_g, 4 emb_ln_b, 8 Wp, 9 bp, 10 Wc, 11 bc,
//      layer l: 16 + 16*l + {0 Wqkv, 1 bqkv, 2 Wo, 3 bo, 4 ln1_g, 5 ln1_b, 6 W1, 7 b1, 8 W2, 9 b2,
//                            10 ln2_g, 11 ln2_b}
std::vector<TensorSpec> tensor_list(const ModelHeader& m) {
  const int64_t H = m.hidden, I = m.intermediate;
  std::vector<TensorSpec> t;
  t.push_back({0, int64_t(m.vocab_size) * H, ROLE_WEIGHT});
  t.push_back({1, int64_t(m.max_positions) * H, ROLE_WEIGHT});
  t.push_back({2, int64_t(m.type_vocab) * H, ROLE_WEIGHT});
  t.push_back({3, H, ROLE_GAMMA});
  t.push_back({4, H, ROLE_BETA});
  for (uint32_t l = 0; l < m.layers; ++l) {
    const uint32_t b = 16 + 16 * l;
    t.push_back({b + 0, 3 * H * H, ROLE_WEIGHT});
    t.push_back({b + 1, 3 * H, ROLE_BIAS});
    t.push_back({b + 2, H * H, ROLE_WEIGHT});
    t.push_back({b + 3, H, ROLE_BIAS});
    t.push_back({b + 4, H, ROLE_GAMMA});
    t.push_back({b + 5, H, ROLE_BETA});
    t.push_back({b + 6, I * H, ROLE_WEIGHT});
    t.push_back({b + 7, I, ROLE_BIAS});
    t.push_back({b + 8, H * I, ROLE_WEIGHT});
    t.push_back({b + 9, H, ROLE_BIAS});
    t.push_back({b + 10, H, ROLE_GAMMA});
    t.push_back({b + 11, H, ROLE_BETA});
  }
  t.push_back({8, H * H, ROLE_WEIGHT});
  t.push_back({9, H, ROLE_BIAS});
  t.push_back({10, int64_t(m.num_labels) * H, ROLE_CLASSIFIER});
  t.push_back({11, int64_t(m.num_labels), ROLE_BIAS});
  return t;
}

struct LayerWeights {
  __nv_bfloat16 *Wqkv, *Wo, *W1, *W2;
  float *bqkv, *bo, *b1, *b2, *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  CUtensorMap tm_qkv, tm_o, tm_1, tm_2;      // W maps for the one-CTA GEMM (256-row boxes)
  CUtensorMap tm2_qkv, tm2_o, tm2_1, tm2_2;  // W maps for the CTA-pair GEMM (128-row boxes)
  // LayerNorm folded into the Linear that follows it (GemmLnArgs in b2_kernels.h): the QKV
  // projection absorbs the previous layer's output LayerNorm, FFN1 absorbs this layer's attention
  // LayerNorm. W*_f = bf16(gamma ∘ W), g = its row sums, c = bias + W·beta.
  __nv_bfloat16 *Wqkv_f = nullptr, *W1_f = nullptr;
  float *g_qkv = nullptr, *c_qkv = nullptr, *g_1 = nullptr, *c_1 = nullptr;
  CUtensorMap tm2_qkv_f, tm2_1_f;
};

enum Kind {
  K_TOKENIZE, K_EMBED, K_GEMM_QKV, K_ATTENTION, K_GEMM_OUT, K_LAYERNORM, K_GEMM_FFN1, K_GEMM_FFN2,
  K_POOLER, K_HEAD, K_COUNT
};

}  // namespace
}  // namespace b2

using namespace b2;

// Everything one batch in flight needs (rows = max_batch * max_len).
struct BatchBuffers {
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;  // recorded behind the lane's last batch of a tokenizer run / job
  int32_t *ids = nullptr, *lens = nullptr;
  uint8_t* mask = nullptr;
  __nv_bfloat16 *x = nullptr, *qkv = nullptr, *ctx = nullptr, *y = nullptr, *ffn = nullptr,
                *pooled = nullptr;
  CUtensorMap tm_x{}, tm_ctx{}, tm_y{}, tm_ffn{};  // A operands; also TMA store / residual maps
  CUtensorMap tm_qkv_c{};                          // qkv as a GEMM output
  AttentionMaps tm_att{};
  // packed (padding-free) layout: token rows of a batch are stored back to back, sequence b at
  // rows [cu[b], cu[b+1]); cu[B] is the batch's row count, read by the kernels from device memory
  int32_t* cu = nullptr;
  // folded LayerNorm: (mean, rstd) per row and the partial sums they are made from
  float2 *stats_a = nullptr, *stats_b = nullptr, *part = nullptr, *part_b = nullptr;
  // [CLS]-only tail of the last layer: after its attention only row 0 of every sequence feeds the
  // pooler, so the out-projection, FFN and both LayerNorms of that layer run on [B, *] blocks
  __nv_bfloat16 *cls = nullptr, *cls_ctx = nullptr, *cls_y = nullptr, *cls_f = nullptr;
  CUtensorMap tm_cls{}, tm_cls_ctx{}, tm_cls_y{}, tm_cls_f{};
};

struct b2_handle {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  std::mutex mu;

  bool has_tokenizer = false;
  TokenizerDev tok{};
  int tok_vocab = 0;

  bool has_model = false;
  ModelHeader mh{};
  float *word = nullptr, *pos = nullptr, *type0 = nullptr, *emb_g = nullptr, *emb_b = nullptr;
  std::vector<LayerWeights> layers;
  __nv_bfloat16* Wp = nullptr;
  float *bp = nullptr, *Wc = nullptr, *bc = nullptr;
  CUtensorMap tm_p{};
  int64_t weight_bytes = 0, activation_bytes = 0;
  std::vector<void*> allocations;

  int max_batch = 1024, max_len = 128;
  // Two sets of per-batch buffers, each with its own stream: consecutive batches of a job alternate
  // between them, so the ragged end of one batch's kernels (a last wave that fills 27% of the SMs,
  // the few microseconds every launch needs to ramp up and drain, the small bandwidth-bound kernels
  // between the GEMMs) is covered by the other batch's work. lane[0] runs on `stream`.
  BatchBuffers lane[2];
  int n_lanes = 1;
  unsigned long long* hist = nullptr;
  int64_t* d_tokens = nullptr;  // running sum of lens
  int32_t* d_flag = nullptr;    // lowest row of an ids job whose mask is not a non-empty prefix
  // developer aid (b2_debug_hidden_states): when set, forward_batch leaves an f32 copy of every
  // layer's output in dbg_hidden[layer][b][s][:] (one batch only)
  float* dbg_hidden = nullptr;
  bool packed = true;      // padding-free token rows (B2_FORCE_PADDED=1 turns it off)
  bool pair_gemm = true;   // CTA-pair GEMM for the encoder projections (B2_GEMM_1CTA=1 turns it off)
  // LayerNorm folding (B2_NO_LN_FOLD=1 turns it off): no LayerNorm kernel runs in the encoder; the
  // un-normalised sums stay in x / y and the row statistics travel beside them
  bool fold_ln = true;
  float *ones = nullptr, *zeros = nullptr;
  int part_slots = 0;

  // token ids / lengths of a run of items of the job: the tokenizer runs once per TOKENIZE_RUN items
  // (one launch fills the GPU; one launch per 1,024-item batch leaves most of it idle)
  int32_t* job_ids = nullptr; size_t job_ids_cap = 0;
  int32_t* job_lens = nullptr; size_t job_lens_cap = 0;
  // pinned staging and side streams for large host-to-device copies (b2_csv_open), made on first use
  static constexpr int COPY_LANES = 4;
  uint8_t* pin_buf[COPY_LANES][2] = {};
  cudaEvent_t pin_ev[COPY_LANES][2] = {};
  cudaStream_t copy_stream[COPY_LANES] = {};
  // growable staging for host-pointer calls
  uint8_t* st_text = nullptr; size_t st_text_cap = 0;
  int64_t* st_off = nullptr; size_t st_off_cap = 0;
  void* st_out = nullptr; size_t st_out_cap = 0;

  // profiling
  bool profiling = false;
  std::vector<cudaEvent_t> events;
  std::vector<int> event_kinds;
  size_t n_events = 0;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  cudaEvent_t ev_run = nullptr;  // inputs of the current tokenizer run / job are ready (lane 1 waits on it)
};

namespace b2 {
namespace {

int dev_alloc(b2_handle* h, void** p, size_t bytes) {
  B2_CUDA_OK(cudaMalloc(p, bytes ? bytes : 16));
  h->allocations.push_back(*p);
  return B2_OK;
}

int grow(void** p, size_t* cap, size_t need) {
  if (need <= *cap) return B2_OK;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  size_t want = need + need / 4 + 4096;
  B2_CUDA_OK(cudaMalloc(p, want));
  *cap = want;
  return B2_OK;
}

float synth_scale(float std_dev) {
  // std of the sum of four independent uniform integers in [0, 65535]
  const double sd = sqrt(4.0 * (65536.0 * 65536.0 - 1.0) / 12.0);
  return static_cast<float>(static_cast<double>(std_dev) / sd);
}

// mark: profiling boundary before a launch of kind `k`
void mark(b2_handle* h, int kind) {
  if (!h->profiling) return;
  if (h->n_events == h->events.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    h->events.push_back(e);
    h->event_kinds.push_back(0);
  }
  h->event_kinds[h->n_events] = kind;
  cudaEventRecord(h->events[h->n_events], h->stream);
  h->n_events++;
}

struct KindAcc {
  double ms[K_COUNT] = {};
  int64_t n[K_COUNT] = {};
  int64_t launches = 0;
};

// close the current profiling segment list: events[i] .. events[i+1] brackets a launch of kind[i]
int flush_profile(b2_handle* h, KindAcc* acc) {
  if (!h->profiling || h->n_events == 0) return B2_OK;
  mark(h, -1);
  B2_CUDA_OK(cudaStreamSynchronize(h->stream));
  for (size_t i = 0; i + 1 < h->n_events; ++i) {
    float ms = 0.f;
    B2_CUDA_OK(cudaEventElapsedTime(&ms, h->events[i], h->events[i + 1]));
    const int k = h->event_kinds[i];
    if (k >= 0 && k < K_COUNT) { acc->ms[k] += ms; acc->n[k] += 1; }
  }
  h->n_events = 0;
  return B2_OK;
}

int load_model(b2_handle* h, const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) {
    set_last_error("model file not found: %s", path);
    return B2_ERR_IO;
  }
  ModelHeader& m = h->mh;
  if (fread(&m, sizeof(m), 1, f) != 1 || memcmp(m.magic, "B2MD", 4) != 0 || m.version != 1) {
    fclose(f);
    set_last_error("'%s' is not a B2MD v1 model file", path);
    return B2_ERR_FORMAT;
  }
  if (m.hidden == 0 || m.hidden % 128 != 0 || m.hidden > 1024 || m.heads * 64 != m.hidden ||
      m.intermediate == 0 || m.intermediate % 8 != 0 || m.layers == 0 || m.num_labels == 0 ||
      m.num_labels > 2048 || m.type_vocab == 0 || m.vocab_size == 0 || m.max_positions == 0 || m.kind > 1) {
    fclose(f);
    set_last_error("model '%s': unsupported geometry (hidden=%u must be a multiple of 128 and <= 1024 "
                   "with head dim 64; heads=%u intermediate=%u labels=%u)",
                   path, m.hidden, m.heads, m.intermediate, m.num_labels);
    return B2_ERR_FORMAT;
  }
  const std::vector<TensorSpec> specs = tensor_list(m);
  int64_t max_n = 0;
  for (const auto& s : specs) max_n = std::max(max_n, s.n);
  float* scratch = nullptr;  // f32 staging for tensors that are stored as bf16
  {
    const cudaError_t e = cudaMalloc(&scratch, max_n * sizeof(float));
    if (e != cudaSuccess) {
      fclose(f);
      set_last_error("model '%s': cannot allocate %lld bytes of staging: %s", path,
                     static_cast<long long>(max_n * sizeof(float)), cudaGetErrorString(e));
      return B2_ERR_CUDA;
    }
  }
  std::vector<float> host;
  h->layers.resize(m.layers);
  int rc = B2_OK;

  auto materialise = [&](const TensorSpec& s, float* dst) -> int {
    if (m.kind == 1) {
      float sd = 0.f, shift = 0.f;
      switch (s.role) {
        case ROLE_WEIGHT: sd = m.init_std; break;
        case ROLE_BIAS: sd = m.bias_std; break;
        case ROLE_GAMMA: sd = m.gamma_std; shift = 1.0f; break;
        case ROLE_BETA: sd = m.bias_std; break;
        case ROLE_CLASSIFIER: sd = m.classifier_std; break;
      }
      return fill_synthetic_f32(h->stream, dst, s.n, m.seed, s.id, synth_scale(sd), shift);
    }
    host.resize(s.n);
    if (fread(host.data(), sizeof(float), s.n, f) != static_cast<size_t>(s.n)) {
      set_last_error("model '%s' is truncated (tensor id %u)", path, s.id);
      return B2_ERR_FORMAT;
    }
    B2_CUDA_OK(cudaMemcpyAsync(dst, host.data(), s.n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    B2_CUDA_OK(cudaStreamSynchronize(h->stream));
    return B2_OK;
  };
  auto want_f32 = [&](const TensorSpec& s, float** out) -> int {
    if ((rc = dev_alloc(h, reinterpret_cast<void**>(out), s.n * sizeof(float)))) return rc;
    h->weight_bytes += s.n * sizeof(float);
    return materialise(s, *out);
  };
  auto want_bf16 = [&](const TensorSpec& s, __nv_bfloat16** out) -> int {
    if ((rc = dev_alloc(h, reinterpret_cast<void**>(out), s.n * sizeof(__nv_bfloat16)))) return rc;
    h->weight_bytes += s.n * sizeof(__nv_bfloat16);
    if ((rc = materialise(s, scratch))) return rc;
    return convert_f32_to_bf16(h->stream, scratch, *out, s.n);
  };

  size_t k = 0;
  do {
    if ((rc = want_f32(specs[k++], &h->word))) break;
    if ((rc = want_f32(specs[k++], &h->pos))) break;
    if ((rc = want_f32(specs[k++], &h->type0))) break;
    if ((rc = want_f32(specs[k++], &h->emb_g))) break;
    if ((rc = want_f32(specs[k++], &h->emb_b))) break;
    for (uint32_t l = 0; l < m.layers && !rc; ++l) {
      LayerWeights& L = h->layers[l];
      const int Hh = m.hidden, Ii = m.intermediate;
      if ((rc = want_bf16(specs[k++], &L.Wqkv))) break;   // f32 copy stays in `scratch` until folded
      if ((rc = want_f32(specs[k++], &L.bqkv))) break;
      {
        // the QKV projection reads the previous layer's output LayerNorm (none before layer 0)
        const float* pg = l ? h->layers[l - 1].ln2_g : nullptr;
        const float* pb = l ? h->layers[l - 1].ln2_b : nullptr;
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.Wqkv_f), size_t(3) * Hh * Hh * 2))) break;
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.g_qkv), size_t(3) * Hh * 4))) break;
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.c_qkv), size_t(3) * Hh * 4))) break;
        h->weight_bytes += size_t(3) * Hh * Hh * 2;
        if ((rc = fold_ln_into_linear(h->stream, scratch, L.bqkv, pg, pb, 3 * Hh, Hh, L.Wqkv_f, L.g_qkv,
                                      L.c_qkv))) break;
      }
      if ((rc = want_bf16(specs[k++], &L.Wo))) break;
      if ((rc = want_f32(specs[k++], &L.bo))) break;
      if ((rc = want_f32(specs[k++], &L.ln1_g))) break;
      if ((rc = want_f32(specs[k++], &L.ln1_b))) break;
      if ((rc = want_bf16(specs[k++], &L.W1))) break;
      if ((rc = want_f32(specs[k++], &L.b1))) break;
      {
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.W1_f), size_t(Ii) * Hh * 2))) break;
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.g_1), size_t(Ii) * 4))) break;
        if ((rc = dev_alloc(h, reinterpret_cast<void**>(&L.c_1), size_t(Ii) * 4))) break;
        h->weight_bytes += size_t(Ii) * Hh * 2;
        if ((rc = fold_ln_into_linear(h->stream, scratch, L.b1, L.ln1_g, L.ln1_b, Ii, Hh, L.W1_f, L.g_1,
                                      L.c_1))) break;
      }
      if ((rc = want_bf16(specs[k++], &L.W2))) break;
      if ((rc = want_f32(specs[k++], &L.b2))) break;
      if ((rc = want_f32(specs[k++], &L.ln2_g))) break;
      if ((rc = want_f32(specs[k++], &L.ln2_b))) break;
      const int H = m.hidden, I = m.intermediate;
      if ((rc = gemm_tmap_b(&L.tm_qkv, L.Wqkv, 3 * H, H))) break;
      if ((rc = gemm_tmap_b(&L.tm_o, L.Wo, H, H))) break;
      if ((rc = gemm_tmap_b(&L.tm_1, L.W1, I, H))) break;
      if ((rc = gemm_tmap_b(&L.tm_2, L.W2, H, I))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_qkv, L.Wqkv, 3 * H, H))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_o, L.Wo, H, H))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_1, L.W1, I, H))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_2, L.W2, H, I))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_qkv_f, L.Wqkv_f, 3 * H, H))) break;
      if ((rc = gemm2_tmap_b(&L.tm2_1_f, L.W1_f, I, H))) break;
    }
    if (rc) break;
    if ((rc = want_bf16(specs[k++], &h->Wp))) break;
    if ((rc = want_f32(specs[k++], &h->bp))) break;
    if ((rc = want_f32(specs[k++], &h->Wc))) break;
    if ((rc = want_f32(specs[k++], &h->bc))) break;
    if ((rc = gemm_tmap_b(&h->tm_p, h->Wp, m.hidden, m.hidden))) break;
  } while (0);
  fclose(f);
  cudaError_t e = cudaStreamSynchronize(h->stream);
  cudaFree(scratch);
  if (rc) return rc;
  if (e != cudaSuccess) {
    set_last_error("model load failed on device: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  h->has_model = true;
  return B2_OK;
}

int alloc_lane(b2_handle* h, BatchBuffers& L) {
  const ModelHeader& m = h->mh;
  const int64_t rows = int64_t(h->max_batch) * h->max_len;
  const int64_t H = m.hidden, I = m.intermediate;
  int rc;
  auto A = [&](void** p, int64_t bytes) { h->activation_bytes += bytes; return dev_alloc(h, p, bytes); };
  if ((rc = A(reinterpret_cast<void**>(&L.ids), rows * 4))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.mask), rows))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.lens), h->max_batch * 4))) return rc;
  if (!h->has_model) return B2_OK;
  B2_CUDA_OK(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
  if ((rc = A(reinterpret_cast<void**>(&L.x), rows * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.qkv), rows * 3 * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.ctx), rows * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.y), rows * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.ffn), rows * I * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.pooled), int64_t(h->max_batch) * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.cu), (int64_t(h->max_batch) + 1) * 4))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.stats_a), rows * 8))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.stats_b), rows * 8))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.part), rows * h->part_slots * 8))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.part_b), rows * h->part_slots * 8))) return rc;
  B2_CUDA_OK(cudaMemsetAsync(L.part, 0, rows * h->part_slots * 8, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.part_b, 0, rows * h->part_slots * 8, h->stream));
  if ((rc = A(reinterpret_cast<void**>(&L.cls), int64_t(h->max_batch) * H * 2))) return rc;
  B2_CUDA_OK(cudaMemsetAsync(L.cls, 0, int64_t(h->max_batch) * H * 2, h->stream));
  if ((rc = gemm_tmap_a(&L.tm_cls, L.cls, h->max_batch, H, H))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.cls_ctx), int64_t(h->max_batch) * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.cls_y), int64_t(h->max_batch) * H * 2))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&L.cls_f), int64_t(h->max_batch) * I * 2))) return rc;
  B2_CUDA_OK(cudaMemsetAsync(L.cls_ctx, 0, int64_t(h->max_batch) * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.cls_y, 0, int64_t(h->max_batch) * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.cls_f, 0, int64_t(h->max_batch) * I * 2, h->stream));
  if ((rc = gemm_tmap_a(&L.tm_cls_ctx, L.cls_ctx, h->max_batch, H, H))) return rc;
  if ((rc = gemm_tmap_a(&L.tm_cls_y, L.cls_y, h->max_batch, H, H))) return rc;
  if ((rc = gemm_tmap_a(&L.tm_cls_f, L.cls_f, h->max_batch, I, I))) return rc;
  // pad rows are read by TMA boxes and attention tiles before they are ever written: give them
  // finite contents once so no NaN bit pattern can reach a masked-out product
  B2_CUDA_OK(cudaMemsetAsync(L.x, 0, rows * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.qkv, 0, rows * 3 * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.ctx, 0, rows * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.y, 0, rows * H * 2, h->stream));
  B2_CUDA_OK(cudaMemsetAsync(L.ffn, 0, rows * I * 2, h->stream));
  if ((rc = gemm_tmap_a(&L.tm_x, L.x, rows, H, H))) return rc;
  if ((rc = gemm_tmap_a(&L.tm_ctx, L.ctx, rows, H, H))) return rc;
  if ((rc = gemm_tmap_a(&L.tm_y, L.y, rows, H, H))) return rc;
  if ((rc = gemm_tmap_a(&L.tm_ffn, L.ffn, rows, I, I))) return rc;
  if ((rc = gemm2_tmap_c(&L.tm_qkv_c, L.qkv, rows, 3 * H, 3 * H))) return rc;
  if ((rc = attention_tmaps(&L.tm_att, L.qkv, rows, H))) return rc;
  return B2_OK;
}

int alloc_activations(b2_handle* h) {
  const ModelHeader& m = h->mh;
  const int64_t H = m.hidden;
  int rc;
  auto A = [&](void** p, int64_t bytes) { h->activation_bytes += bytes; return dev_alloc(h, p, bytes); };
  if ((rc = A(reinterpret_cast<void**>(&h->d_tokens), 8))) return rc;
  if ((rc = A(reinterpret_cast<void**>(&h->d_flag), 8))) return rc;
  h->lane[0].stream = h->stream;
  h->n_lanes = 1;
  if (h->has_model) {
    if ((rc = A(reinterpret_cast<void**>(&h->hist), int64_t(m.num_labels) * 8))) return rc;
    h->part_slots = static_cast<int>((H + 63) / 64);
    if ((rc = A(reinterpret_cast<void**>(&h->ones), H * 4))) return rc;
    if ((rc = A(reinterpret_cast<void**>(&h->zeros), H * 4))) return rc;
    B2_CUDA_OK(cudaMemsetAsync(h->zeros, 0, H * 4, h->stream));
    if ((rc = fill_synthetic_f32(h->stream, h->ones, H, 0, 0, 0.0f, 1.0f))) return rc;  // 0*x + 1
    const char* one = getenv("B2_ONE_LANE");
    if (!(one && one[0] == '1')) {
      B2_CUDA_OK(cudaStreamCreateWithFlags(&h->lane[1].stream, cudaStreamNonBlocking));
      h->n_lanes = 2;
    }
  }
  for (int l = 0; l < h->n_lanes; ++l)
    if ((rc = alloc_lane(h, h->lane[l]))) return rc;
  B2_CUDA_OK(cudaStreamSynchronize(h->stream));
  return B2_OK;
}

constexpr int64_t TOKENIZE_RUN = 32768;  // items per tokenizer launch inside a job

__global__ void sum_lens_kernel(const int32_t* lens, int n, int64_t* total) {
  long long s = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += lens[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ long long part[32];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t = 0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t += part[w];
    atomicAdd(reinterpret_cast<unsigned long long*>(total), static_cast<unsigned long long>(t));  // two lanes
  }
}

__global__ void hist_to_i64_kernel(const unsigned long long* src, int64_t* dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = static_cast<int64_t>(src[i]);
}

// Encoder + head for one batch: ids [B, S] and lens [B] in device memory.
int forward_batch(b2_handle* h, BatchBuffers& bb, const int32_t* ids, const int32_t* lens, int B, int S, int topk,
                  int32_t* d_top_idx, float* d_top_score, int32_t* d_labels, float* d_logits,
                  int64_t* launches) {
  const ModelHeader& m = h->mh;
  const int H = m.hidden, I = m.intermediate, M = B * S;
  cudaStream_t st = bb.stream;
  int rc;
  // packed layout: M is only the upper bound; the kernels read the true row count from cu[B]
  const int32_t* cu = h->packed ? bb.cu : nullptr;
  const int32_t* m_dev = h->packed ? bb.cu + B : nullptr;
  mark(h, K_EMBED);
  if (h->packed) {
    if ((rc = pack_offsets(st, lens, B, bb.cu))) return rc;
    ++*launches;
  }
  if ((rc = embed_layernorm(st, ids, h->word, h->pos, h->type0, h->emb_g, h->emb_b, m.ln_eps, bb.x,
                            M, S, H, m.vocab_size, h->packed ? lens : nullptr, cu))) return rc;
  ++*launches;
  const size_t dbg_layer = static_cast<size_t>(B) * S * H;  // floats per layer in dbg_hidden
  if (h->dbg_hidden &&
      (rc = dump_hidden(st, bb.x, nullptr, nullptr, 0, m.ln_eps, nullptr, nullptr, lens, cu, B, S, H,
                        h->dbg_hidden))) return rc;
  // encoder projections: CTA-pair kernel when the batch is large enough to fill 74 pairs
  const bool pair = h->pair_gemm && M >= 4096;
  auto gemm = [&](const CUtensorMap& ta, const CUtensorMap& tb1, const CUtensorMap& tb2, const CUtensorMap& tc,
                  __nv_bfloat16* Cc, const float* bias, const CUtensorMap* tr, const __nv_bfloat16* res, int N,
                  int K, int ldc, int ldr, int epi) {
    return pair ? gemm2_bf16_tn(st, ta, tb2, tc, tr, bias, M, N, K, epi, h->num_sms, m_dev)
                : gemm_bf16_tn(st, ta, tb1, Cc, bias, res, M, N, K, ldc, ldr, epi, h->num_sms, m_dev);
  };
  const bool fold = h->fold_ln;
  // folded path: x / y hold un-normalised sums; the LayerNorm that turns them into layer inputs is
  // described by (stats, gamma, beta) and applied inside the consuming GEMMs' epilogues
  // Row statistics reach the consuming epilogues as the 12 partial (Σ, Σ²) slots the producing GEMM
  // wrote (in_part; summed in slot order by every reader) — except for the embedding output, which
  // is already normalised (identity statistics in stats_b).
  const float2* in_stats = bb.stats_b;
  const float2* in_part = nullptr;
  const float inv_h = 1.0f / static_cast<float>(H);
  const float* in_g = h->ones;
  const float* in_b = h->zeros;
  if (fold) {
    if ((rc = ln_stats_identity(st, bb.stats_b, M))) return rc;  // the embedding output is normalised
    ++*launches;
  }
  auto gemm_ln = [&](const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const float* bias,
                     const CUtensorMap* tr, int N, int K, int epi, const GemmLnArgs& a) {
    return gemm2_bf16_tn(st, ta, tb, tc, tr, bias, M, N, K, epi, h->num_sms, m_dev, &a);
  };
  for (uint32_t l = 0; l < m.layers; ++l) {
    LayerWeights& L = h->layers[l];
    mark(h, K_GEMM_QKV);
    if (fold) {
      GemmLnArgs a;
      a.row_stats = in_stats; a.part_in = in_part; a.part_slots = h->part_slots; a.inv_h = inv_h; a.eps = m.ln_eps;
      a.g = L.g_qkv;
      if ((rc = gemm_ln(bb.tm_x, L.tm2_qkv_f, bb.tm_qkv_c, L.c_qkv, nullptr, 3 * H, H, GEMM_EPI_LNIN, a)))
        return rc;
    } else if ((rc = gemm(bb.tm_x, L.tm_qkv, L.tm2_qkv, bb.tm_qkv_c, bb.qkv, L.bqkv, nullptr, nullptr, 3 * H, H,
                          3 * H, 0, GEMM_EPI_BIAS))) {
      return rc;
    }
    mark(h, K_ATTENTION);
    if ((rc = attention_bf16(st, bb.tm_att, bb.ctx, lens, B, S, H, cu))) return rc;
    if (l + 1 == m.layers) {
      // last layer: only the [CLS] rows are read downstream (pooler), and every remaining kernel is
      // row-wise, so run them on the B gathered rows; per-row results are unchanged
      mark(h, K_GEMM_OUT);
      if ((rc = gather_cls_rows(st, bb.ctx, cu, S, B, H, bb.cls_ctx))) return rc;
      if (fold) {
        if (in_part) {  // the gather wants finished (mean, rstd): one finalize per batch, not one per GEMM
          if ((rc = ln_stats_finalize(st, in_part, h->part_slots, M, H, m.ln_eps, bb.stats_a, m_dev))) return rc;
          in_stats = bb.stats_a;
          ++*launches;
        }
        if ((rc = gather_cls_rows_norm(st, bb.x, in_stats, in_g, in_b, cu, S, B, H, bb.cls))) return rc;
      } else if ((rc = gather_cls_rows(st, bb.x, cu, S, B, H, bb.cls))) {
        return rc;
      }
      if ((rc = gemm_bf16_tn(st, bb.tm_cls_ctx, L.tm_o, bb.cls_y, L.bo, bb.cls, B, H, H, H, H,
                             GEMM_EPI_BIAS_RESIDUAL, h->num_sms))) return rc;
      mark(h, K_LAYERNORM);
      if ((rc = layernorm_inplace(st, bb.cls_y, L.ln1_g, L.ln1_b, m.ln_eps, B, H))) return rc;
      mark(h, K_GEMM_FFN1);
      if ((rc = gemm_bf16_tn(st, bb.tm_cls_y, L.tm_1, bb.cls_f, L.b1, nullptr, B, I, H, I, 0,
                             GEMM_EPI_BIAS_GELU, h->num_sms))) return rc;
      mark(h, K_GEMM_FFN2);
      if ((rc = gemm_bf16_tn(st, bb.tm_cls_f, L.tm_2, bb.cls, L.b2, bb.cls_y, B, H, I, H, H,
                             GEMM_EPI_BIAS_RESIDUAL, h->num_sms))) return rc;
      mark(h, K_LAYERNORM);
      if ((rc = layernorm_inplace(st, bb.cls, L.ln2_g, L.ln2_b, m.ln_eps, B, H))) return rc;
      *launches += 9;
      if (h->dbg_hidden && (rc = dump_cls(st, bb.cls, B, S, H, h->dbg_hidden + (l + 1) * dbg_layer))) return rc;
      break;
    }
    if (fold) {
      GemmLnArgs a;
      mark(h, K_GEMM_OUT);
      a.row_stats = in_stats; a.part_in = in_part; a.inv_h = inv_h; a.eps = m.ln_eps;
      a.res_gamma = in_g; a.res_beta = in_b;
      a.part_out = bb.part; a.part_slots = h->part_slots;  // statistics of y
      if ((rc = gemm_ln(bb.tm_ctx, L.tm2_o, bb.tm_y, L.bo, &bb.tm_x, H, H, GEMM_EPI_LNRES_STATS, a))) return rc;
      // y's statistics feed the two widest epilogues (FFN1 has 12 column tiles per row block, each
      // of which would re-add the 12 slots): finish them once; x's statistics (below) are read as
      // partials by QKV and the out-projection, which saves the second finalize of the layer
      mark(h, K_LAYERNORM);
      if ((rc = ln_stats_finalize(st, bb.part, h->part_slots, M, H, m.ln_eps, bb.stats_a, m_dev))) return rc;
      mark(h, K_GEMM_FFN1);
      GemmLnArgs f1;
      f1.row_stats = bb.stats_a;
      f1.g = L.g_1;
      if ((rc = gemm_ln(bb.tm_y, L.tm2_1_f, bb.tm_ffn, L.c_1, nullptr, I, H, GEMM_EPI_LNIN_GELU, f1)))
        return rc;
      mark(h, K_GEMM_FFN2);
      GemmLnArgs f2;
      f2.row_stats = bb.stats_a;
      f2.res_gamma = L.ln1_g; f2.res_beta = L.ln1_b;
      f2.part_out = bb.part_b; f2.part_slots = h->part_slots;  // statistics of the next layer's x
      if ((rc = gemm_ln(bb.tm_ffn, L.tm2_2, bb.tm_x, L.b2, &bb.tm_y, H, I, GEMM_EPI_LNRES_STATS, f2))) return rc;
      in_part = bb.part_b; in_g = L.ln2_g; in_b = L.ln2_b;
      *launches += 6;
      if (h->dbg_hidden &&
          (rc = dump_hidden(st, bb.x, nullptr, in_part, h->part_slots, m.ln_eps, in_g, in_b, lens, cu, B, S, H,
                            h->dbg_hidden + (l + 1) * dbg_layer))) return rc;
      continue;
    }
    mark(h, K_GEMM_OUT);
    if ((rc = gemm(bb.tm_ctx, L.tm_o, L.tm2_o, bb.tm_y, bb.y, L.bo, &bb.tm_x, bb.x, H, H, H, H,
                   GEMM_EPI_BIAS_RESIDUAL)))
      return rc;
    mark(h, K_LAYERNORM);
    if ((rc = layernorm_inplace(st, bb.y, L.ln1_g, L.ln1_b, m.ln_eps, M, H, m_dev))) return rc;
    mark(h, K_GEMM_FFN1);
    if ((rc = gemm(bb.tm_y, L.tm_1, L.tm2_1, bb.tm_ffn, bb.ffn, L.b1, nullptr, nullptr, I, H, I, 0,
                   GEMM_EPI_BIAS_GELU)))
      return rc;
    mark(h, K_GEMM_FFN2);
    if ((rc = gemm(bb.tm_ffn, L.tm_2, L.tm2_2, bb.tm_x, bb.x, L.b2, &bb.tm_y, bb.y, H, I, H, H,
                   GEMM_EPI_BIAS_RESIDUAL)))
      return rc;
    mark(h, K_LAYERNORM);
    if ((rc = layernorm_inplace(st, bb.x, L.ln2_g, L.ln2_b, m.ln_eps, M, H, m_dev))) return rc;
    *launches += 7;
    if (h->dbg_hidden &&
        (rc = dump_hidden(st, bb.x, nullptr, nullptr, 0, m.ln_eps, nullptr, nullptr, lens, cu, B, S, H,
                          h->dbg_hidden + (l + 1) * dbg_layer))) return rc;
  }
  mark(h, K_POOLER);
  if ((rc = gemm_bf16_tn(st, bb.tm_cls, h->tm_p, bb.pooled, h->bp, nullptr, B, H, H, H, 0,
                         GEMM_EPI_BIAS_TANH, h->num_sms))) return rc;
  mark(h, K_HEAD);
  if ((rc = classify_head(st, bb.pooled, H, h->Wc, h->bc, B, H, m.num_labels, topk, d_logits,
                          d_top_idx, d_top_score, d_labels, h->hist))) return rc;
  *launches += 2;
  return B2_OK;
}

void fill_metrics(b2_metrics* out, const b2_handle* h, const KindAcc& acc, int64_t items,
                  int64_t tokens, int64_t padded, int64_t batches, int64_t h2d, int64_t d2h,
                  double gpu_ms) {
  if (!out) return;
  b2_metrics mtr;
  memset(&mtr, 0, sizeof(mtr));
  mtr.struct_size = sizeof(b2_metrics);
  mtr.profiled = h->profiling ? 1 : 0;
  mtr.items = items; mtr.tokens = tokens; mtr.padded_tokens = padded; mtr.batches = batches;
  mtr.kernel_launches = acc.launches; mtr.h2d_bytes = h2d; mtr.d2h_bytes = d2h;
  mtr.gpu_ms_total = gpu_ms;
  mtr.ms_tokenize = acc.ms[K_TOKENIZE]; mtr.n_tokenize = acc.n[K_TOKENIZE];
  mtr.ms_embed = acc.ms[K_EMBED]; mtr.n_embed = acc.n[K_EMBED];
  mtr.ms_gemm_qkv = acc.ms[K_GEMM_QKV]; mtr.n_gemm_qkv = acc.n[K_GEMM_QKV];
  mtr.ms_attention = acc.ms[K_ATTENTION]; mtr.n_attention = acc.n[K_ATTENTION];
  mtr.ms_gemm_attn_out = acc.ms[K_GEMM_OUT]; mtr.n_gemm_attn_out = acc.n[K_GEMM_OUT];
  mtr.ms_layernorm = acc.ms[K_LAYERNORM]; mtr.n_layernorm = acc.n[K_LAYERNORM];
  mtr.ms_gemm_ffn1 = acc.ms[K_GEMM_FFN1]; mtr.n_gemm_ffn1 = acc.n[K_GEMM_FFN1];
  mtr.ms_gemm_ffn2 = acc.ms[K_GEMM_FFN2]; mtr.n_gemm_ffn2 = acc.n[K_GEMM_FFN2];
  mtr.ms_pooler = acc.ms[K_POOLER]; mtr.n_pooler = acc.n[K_POOLER];
  mtr.ms_head = acc.ms[K_HEAD]; mtr.n_head = acc.n[K_HEAD];
  const size_t n = std::min<size_t>(sizeof(mtr), out->struct_size > 0 ? out->struct_size : 0);
  const int32_t keep = out->struct_size;
  memcpy(out, &mtr, n);
  out->struct_size = keep;
}

int check_len(const b2_handle* h, int32_t max_len, bool need_model) {
  if (max_len < 2 || max_len > h->max_len) {
    set_last_error("max_len %d is outside [2, %d] (the handle was created for max_len %d)", max_len,
                   h->max_len, h->max_len);
    return B2_ERR_LIMIT;
  }
  if (need_model) {
    if (!h->has_model) {
      set_last_error("this handle was created without a model (b2_config.model_path was NULL)");
      return B2_ERR_ARG;
    }
    if (max_len % 128 != 0) {
      set_last_error("classification needs max_len to be a multiple of 128 (got %d)", max_len);
      return B2_ERR_ARG;
    }
    if (static_cast<uint32_t>(max_len) > h->mh.max_positions) {
      set_last_error("max_len %d exceeds the model's %u positions", max_len, h->mh.max_positions);
      return B2_ERR_LIMIT;
    }
  }
  return B2_OK;
}

// Core of classification with everything in device memory. If d_utf8 is null, ids / mask rows are
// taken from d_ids / d_mask instead of the tokenizer.
int classify_device(b2_handle* h, const uint8_t* d_utf8, const int64_t* d_offsets,
                    const int32_t* d_ids_in, const uint8_t* d_mask_in, int64_t n, int32_t max_len,
                    int32_t topk, int32_t* d_top_idx, float* d_top_score, int32_t* d_labels,
                    int64_t* d_hist, float* d_logits, b2_metrics* metrics, int64_t h2d, int64_t d2h) {
  int rc;
  if ((rc = check_len(h, max_len, true))) return rc;
  const int C = h->mh.num_labels;
  if (topk < 1 || topk > C) {
    set_last_error("topk %d is outside [1, num_labels = %d]", topk, C);
    return B2_ERR_ARG;
  }
  if (n < 0 || (n > 0 && (!d_top_idx || !d_top_score))) {
    set_last_error("classify: n must be >= 0 and top_idx / top_score must not be NULL");
    return B2_ERR_ARG;
  }
  if (d_utf8 && !h->has_tokenizer) {
    set_last_error("this handle was created without a tokenizer (b2_config.vocab_path was NULL)");
    return B2_ERR_ARG;
  }
  cudaStream_t st = h->stream;
  KindAcc acc;
  // a previous job that failed half way may have left work on the second lane's stream
  if (h->n_lanes > 1) B2_CUDA_OK(cudaStreamSynchronize(h->lane[1].stream));
  B2_CUDA_OK(cudaMemsetAsync(h->hist, 0, size_t(C) * 8, st));
  B2_CUDA_OK(cudaMemsetAsync(h->d_tokens, 0, 8, st));
  B2_CUDA_OK(cudaEventRecord(h->ev_begin, st));
  int64_t batches = 0;
  // Text is tokenised TOKENIZE_RUN items at a time, ahead of the batches that consume it.
  const int64_t run_items = std::max<int64_t>(h->max_batch, TOKENIZE_RUN / h->max_batch * h->max_batch);
  if (d_utf8 && n > 0) {
    const int64_t cap = std::min<int64_t>(n, run_items);
    if ((rc = grow(reinterpret_cast<void**>(&h->job_ids), &h->job_ids_cap, size_t(cap) * max_len * 4)))
      return rc;
    if ((rc = grow(reinterpret_cast<void**>(&h->job_lens), &h->job_lens_cap, size_t(cap) * 4))) return rc;
  }
  int64_t run0 = 0;  // first item of the current tokenizer run
  // Batches alternate between the lanes (one lane while profiling: its events live on one stream).
  const int lanes = h->profiling ? 1 : h->n_lanes;
  bool lane1_busy = false;  // lane 1 has work the main stream has not waited for yet
  auto join_lane1 = [&]() -> int {
    if (!lane1_busy) return B2_OK;
    B2_CUDA_OK(cudaEventRecord(h->lane[1].done, h->lane[1].stream));
    B2_CUDA_OK(cudaStreamWaitEvent(st, h->lane[1].done, 0));
    lane1_busy = false;
    return B2_OK;
  };
  if (!d_utf8 && n > 0) {
    // Pre-tokenised rows: lengths for the whole job first, and with them the check that every mask
    // row is a non-empty prefix of ones. The encoder only knows sequences as prefixes [0, len) of
    // their rows, so a mask with a hole, left padding or no token at all would be classified as a
    // different input without any sign of it; such a job is refused before anything runs.
    if ((rc = grow(reinterpret_cast<void**>(&h->job_lens), &h->job_lens_cap, size_t(n) * 4))) return rc;
    B2_CUDA_OK(cudaMemsetAsync(h->d_flag, 0x7f, 4, st));
    if ((rc = mask_to_lens(st, d_mask_in, h->job_lens, static_cast<int>(n), max_len, h->d_flag))) return rc;
    sum_lens_kernel<<<1, 1024, 0, st>>>(h->job_lens, static_cast<int>(n), h->d_tokens);
    acc.launches += 2;
    int32_t bad = 0;
    B2_CUDA_OK(cudaMemcpyAsync(&bad, h->d_flag, 4, cudaMemcpyDeviceToHost, st));
    B2_CUDA_OK(cudaStreamSynchronize(st));
    if (bad >= 0 && bad < n) {
      set_last_error("classify: mask row %d is not a non-empty prefix of ones (1..10..0); masks with holes, "
                     "left padding or no token are not supported", bad);
      return B2_ERR_ARG;
    }
  }
  if (!d_utf8 && lanes > 1) B2_CUDA_OK(cudaEventRecord(h->ev_run, st));  // staged ids, cleared counters
  bool lane1_synced = false;  // lane 1 has waited on ev_run since it was last recorded
  // labels are needed for the histogram even when the caller does not want them
  for (int64_t i0 = 0; i0 < n; i0 += h->max_batch) {
    const int B = static_cast<int>(std::min<int64_t>(h->max_batch, n - i0));
    const int li = lanes > 1 ? static_cast<int>(batches & 1) : 0;
    BatchBuffers& bb = h->lane[li];
    const int32_t *ids = bb.ids, *lens = bb.lens;
    if (d_utf8) {
      if (i0 == 0 || i0 >= run0 + run_items) {
        run0 = i0;
        const int run_n = static_cast<int>(std::min<int64_t>(run_items, n - run0));
        if ((rc = join_lane1())) return rc;  // lane 1 may still be reading the previous run's ids
        mark(h, K_TOKENIZE);
        if ((rc = tokenizer_run(st, h->tok, d_utf8, d_offsets + run0, run_n, max_len, h->job_ids, nullptr,
                                h->job_lens, h->num_sms))) return rc;
        sum_lens_kernel<<<1, 1024, 0, st>>>(h->job_lens, run_n, h->d_tokens);
        acc.launches += 2;
        if (lanes > 1) B2_CUDA_OK(cudaEventRecord(h->ev_run, st));
        lane1_synced = false;
      }
      ids = h->job_ids + (i0 - run0) * max_len;
      lens = h->job_lens + (i0 - run0);
    }
    if (li == 1 && !lane1_synced) {
      B2_CUDA_OK(cudaStreamWaitEvent(bb.stream, h->ev_run, 0));
      lane1_synced = true;
    }
    if (!d_utf8) {
      ids = d_ids_in + i0 * max_len;
      lens = h->job_lens + i0;
    }
    int32_t* lab = d_labels ? d_labels + i0 : reinterpret_cast<int32_t*>(bb.mask);  // scratch
    if ((rc = forward_batch(h, bb, ids, lens, B, max_len, topk, d_top_idx + i0 * topk,
                            d_top_score + i0 * topk, lab, d_logits ? d_logits + i0 * C : nullptr,
                            &acc.launches))) return rc;
    if (li == 1) lane1_busy = true;
    if ((rc = flush_profile(h, &acc))) return rc;
    ++batches;
  }
  if ((rc = join_lane1())) return rc;
  if (d_hist) {
    hist_to_i64_kernel<<<(C + 255) / 256, 256, 0, st>>>(h->hist, d_hist, C);
    acc.launches += 1;
  }
  B2_CUDA_OK(cudaEventRecord(h->ev_end, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  B2_CUDA_OK(cudaGetLastError());
  float ms = 0.f;
  B2_CUDA_OK(cudaEventElapsedTime(&ms, h->ev_begin, h->ev_end));
  int64_t tokens = 0;
  B2_CUDA_OK(cudaMemcpy(&tokens, h->d_tokens, 8, cudaMemcpyDeviceToHost));
  fill_metrics(metrics, h, acc, n, tokens, n * max_len, batches, h2d, d2h, ms);
  return B2_OK;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace
}  // namespace b2

// ================================================================================================
// exported C ABI
// ================================================================================================
extern "C" {

const char* b2_last_error(void) { return b2::last_error(); }

int32_t b2_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int b2_get_device_info(int32_t device, b2_device_info* out) {
  if (!out) { set_last_error("b2_get_device_info: out is NULL"); return B2_ERR_ARG; }
  cudaDeviceProp p;
  B2_CUDA_OK(cudaGetDeviceProperties(&p, device));
  memset(out, 0, sizeof(*out));
  strncpy(out->name, p.name, sizeof(out->name) - 1);
  out->total_memory_bytes = static_cast<int64_t>(p.totalGlobalMem);
  out->sm_count = p.multiProcessorCount;
  out->cc_major = p.major;
  out->cc_minor = p.minor;
  out->l2_bytes = p.l2CacheSize;
  return B2_OK;
}

int b2_get_device_memory(int32_t device, int64_t* free_bytes, int64_t* total_bytes) {
  const int ndev = b2_device_count();
  if (device < 0 || device >= ndev) {
    set_last_error("b2_get_device_memory: device %d is outside [0, %d)", device, ndev);
    return B2_ERR_ARG;
  }
  DeviceGuard guard(device);
  size_t f = 0, t = 0;
  B2_CUDA_OK(cudaMemGetInfo(&f, &t));
  if (free_bytes) *free_bytes = static_cast<int64_t>(f);
  if (total_bytes) *total_bytes = static_cast<int64_t>(t);
  return B2_OK;
}

int b2_get_peer_info(int32_t device, int32_t peer, int32_t* can_access, int32_t* performance_rank,
                     int32_t* native_atomics) {
  const int ndev = b2_device_count();
  if (device < 0 || device >= ndev || peer < 0 || peer >= ndev) {
    set_last_error("b2_get_peer_info: devices %d, %d are outside [0, %d)", device, peer, ndev);
    return B2_ERR_ARG;
  }
  int ca = 0, pr = 0, na = 0;
  if (device != peer) {
    B2_CUDA_OK(cudaDeviceCanAccessPeer(&ca, device, peer));
    if (ca) {
      B2_CUDA_OK(cudaDeviceGetP2PAttribute(&pr, cudaDevP2PAttrPerformanceRank, device, peer));
      B2_CUDA_OK(cudaDeviceGetP2PAttribute(&na, cudaDevP2PAttrNativeAtomicSupported, device, peer));
    }
  }
  if (can_access) *can_access = ca;
  if (performance_rank) *performance_rank = pr;
  if (native_atomics) *native_atomics = na;
  return B2_OK;
}

int b2_create(const b2_config* cfg, b2_handle** out) {
  if (!cfg || !out || cfg->struct_size < static_cast<int32_t>(sizeof(b2_config))) {
    set_last_error("b2_create: config is NULL or has the wrong struct_size");
    return B2_ERR_ARG;
  }
  *out = nullptr;
  const int ndev = b2_device_count();
  if (ndev <= 0) {
    set_last_error("No CUDA device detected (cudaGetDeviceCount returned 0).");
    return B2_ERR_CUDA;
  }
  if (cfg->device < 0 || cfg->device >= ndev) {
    set_last_error("b2_create: device %d is outside [0, %d)", cfg->device, ndev);
    return B2_ERR_ARG;
  }
  if (cfg->vocab_path && !cfg->unicode_path) {
    set_last_error("b2_create: vocab_path needs unicode_path");
    return B2_ERR_ARG;
  }
  DeviceGuard guard(cfg->device);
  cudaDeviceProp prop;
  B2_CUDA_OK(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) {
    set_last_error("device %d is sm_%d%d; this library holds sm_100a code only", cfg->device,
                   prop.major, prop.minor);
    return B2_ERR_CUDA;
  }
  b2_handle* h = new b2_handle();
  h->device = cfg->device;
  h->num_sms = prop.multiProcessorCount;
  h->max_batch = cfg->max_batch > 0 ? cfg->max_batch : 1024;
  h->max_len = cfg->max_len > 0 ? cfg->max_len : 128;
  // B2_FORCE_PADDED=1 keeps every sequence at max_len rows (the layout the packed path is tested
  // against); the default stores only real tokens
  const char* padded_env = getenv("B2_FORCE_PADDED");
  h->packed = !(padded_env && padded_env[0] == '1');
  const char* one_cta_env = getenv("B2_GEMM_1CTA");
  h->pair_gemm = !(one_cta_env && one_cta_env[0] == '1');
  const char* no_fold_env = getenv("B2_NO_LN_FOLD");
  h->fold_ln = !(no_fold_env && no_fold_env[0] == '1');
  if (h->max_batch > 4096) {
    set_last_error("b2_create: max_batch %d exceeds 4096", h->max_batch);
    delete h;
    return B2_ERR_LIMIT;
  }
  int rc = B2_OK;
  do {
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&h->ev_begin) != cudaSuccess || cudaEventCreate(&h->ev_end) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_run, cudaEventDisableTiming) != cudaSuccess) {
      set_last_error("b2_create: cannot create stream / events: %s", cudaGetErrorString(cudaGetLastError()));
      rc = B2_ERR_CUDA;
      break;
    }
    if (cfg->vocab_path) {
      if ((rc = tokenizer_load(cfg->vocab_path, cfg->unicode_path, &h->tok, &h->tok_vocab))) break;
      h->has_tokenizer = true;
    }
    if (cfg->model_path) {
      if ((rc = load_model(h, cfg->model_path))) break;
      if (static_cast<uint32_t>(h->max_len) > h->mh.max_positions) h->max_len = h->mh.max_positions;
    }
    if ((rc = alloc_activations(h))) break;
  } while (0);
  if (rc) {
    b2_destroy(h);
    return rc;
  }
  *out = h;
  return B2_OK;
}

void b2_destroy(b2_handle* h) {
  if (!h) return;
  DeviceGuard guard(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  for (void* p : h->allocations) cudaFree(p);
  if (h->has_tokenizer) tokenizer_free(&h->tok);
  for (int l = 0; l < b2_handle::COPY_LANES; ++l) {
    for (int b = 0; b < 2; ++b) {
      if (h->pin_buf[l][b]) cudaFreeHost(h->pin_buf[l][b]);
      if (h->pin_ev[l][b]) cudaEventDestroy(h->pin_ev[l][b]);
    }
    if (h->copy_stream[l]) cudaStreamDestroy(h->copy_stream[l]);
  }
  if (h->job_ids) cudaFree(h->job_ids);
  if (h->job_lens) cudaFree(h->job_lens);
  if (h->st_text) cudaFree(h->st_text);
  if (h->st_off) cudaFree(h->st_off);
  if (h->st_out) cudaFree(h->st_out);
  for (cudaEvent_t e : h->events) cudaEventDestroy(e);
  if (h->ev_begin) cudaEventDestroy(h->ev_begin);
  if (h->ev_end) cudaEventDestroy(h->ev_end);
  if (h->ev_run) cudaEventDestroy(h->ev_run);
  for (BatchBuffers& L : h->lane)
    if (L.done) cudaEventDestroy(L.done);
  if (h->lane[1].stream) cudaStreamDestroy(h->lane[1].stream);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

int b2_get_model_info(b2_handle* h, b2_model_info* out) {
  if (!h || !out) { set_last_error("b2_get_model_info: NULL argument"); return B2_ERR_ARG; }
  memset(out, 0, sizeof(*out));
  if (h->has_model) {
    out->hidden = h->mh.hidden; out->layers = h->mh.layers; out->heads = h->mh.heads;
    out->intermediate = h->mh.intermediate; out->vocab_size = h->mh.vocab_size;
    out->max_positions = h->mh.max_positions; out->num_labels = h->mh.num_labels;
  }
  out->tokenizer_vocab_size = h->tok_vocab;
  out->max_batch = h->max_batch; out->max_len = h->max_len; out->device = h->device;
  out->sm_count = h->num_sms;
  out->weight_bytes = h->weight_bytes; out->activation_bytes = h->activation_bytes;
  out->packed_rows = h->packed ? 1 : 0;
  out->ln_folded = h->fold_ln ? 1 : 0;
  return B2_OK;
}

int b2_set_profiling(b2_handle* h, int32_t on) {
  if (!h) { set_last_error("b2_set_profiling: NULL handle"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  h->profiling = on != 0;
  h->n_events = 0;
  return B2_OK;
}

uint64_t b2_get_stream(b2_handle* h) {
  return h ? static_cast<uint64_t>(reinterpret_cast<uintptr_t>(h->stream)) : 0;
}

int b2_tokenize_device(b2_handle* h, const uint8_t* d_utf8, const int64_t* d_offsets, int64_t n,
                       int64_t n_bytes, int32_t max_len, int32_t* d_ids, uint8_t* d_mask,
                       b2_metrics* metrics) {
  (void)n_bytes;
  if (!h) { set_last_error("b2_tokenize_device: NULL handle"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  if (!h->has_tokenizer) {
    set_last_error("this handle was created without a tokenizer (b2_config.vocab_path was NULL)");
    return B2_ERR_ARG;
  }
  if (max_len < 2) { set_last_error("tokenize: max_len must be >= 2"); return B2_ERR_ARG; }
  if (n < 0 || n > 0x7fffffff || (n > 0 && !d_ids)) {
    set_last_error("tokenize: n out of range or ids is NULL");
    return B2_ERR_ARG;
  }
  KindAcc acc;
  B2_CUDA_OK(cudaEventRecord(h->ev_begin, h->stream));
  mark(h, K_TOKENIZE);
  int rc = tokenizer_run(h->stream, h->tok, d_utf8, d_offsets, static_cast<int>(n), max_len, d_ids,
                         d_mask, nullptr, h->num_sms);
  if (rc) return rc;
  acc.launches = n > 0 ? 1 : 0;
  if ((rc = flush_profile(h, &acc))) return rc;
  B2_CUDA_OK(cudaEventRecord(h->ev_end, h->stream));
  B2_CUDA_OK(cudaStreamSynchronize(h->stream));
  B2_CUDA_OK(cudaGetLastError());
  float ms = 0.f;
  B2_CUDA_OK(cudaEventElapsedTime(&ms, h->ev_begin, h->ev_end));
  fill_metrics(metrics, h, acc, n, 0, n * max_len, 1, 0, 0, ms);
  return B2_OK;
}

// copy_text = false: size the buffers and upload the offsets only (the caller uploads the text in pieces)
static int stage_text(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                      int64_t* n_bytes, bool copy_text = true) {
  if (!offsets || n < 0 || n > 0x7fffffff || (n > 0 && !utf8 && offsets[n] > 0)) {
    set_last_error("text input: n out of range or NULL utf8 / offsets");
    return B2_ERR_ARG;
  }
  const int64_t bytes = offsets[n];
  if (offsets[0] != 0 || bytes < 0) {
    set_last_error("text input: offsets must start at 0 and be non-decreasing");
    return B2_ERR_ARG;
  }
  int rc;
  if ((rc = grow(reinterpret_cast<void**>(&h->st_text), &h->st_text_cap, bytes + 16))) return rc;
  if ((rc = grow(reinterpret_cast<void**>(&h->st_off), &h->st_off_cap, (n + 1) * 8))) return rc;
  if (bytes && copy_text) B2_CUDA_OK(cudaMemcpyAsync(h->st_text, utf8, bytes, cudaMemcpyHostToDevice, h->stream));
  B2_CUDA_OK(cudaMemcpyAsync(h->st_off, offsets, (n + 1) * 8, cudaMemcpyHostToDevice, h->stream));
  *n_bytes = bytes;
  return B2_OK;
}

// b2_tokenize for large jobs. In one piece, the upload of the text (0.5 GB for a million lines), the
// kernel and the read-back of ids and masks (0.64 GB) take about 20 ms each and run back to back, which
// makes the end-to-end rate half the kernel's. Here the job is cut into chunks of TOKENIZE_PIPE_CHUNK
// lines that go round three side streams, each chunk uploading its own byte range, running the kernel
// on its lines and sending its rows back: chunk k's results travel to the host while chunk k+1 is
// tokenised and chunk k+2's text arrives (the two copy directions have their own engines). The kernel
// reads lines at their absolute offsets, so every chunk lands where a one-shot copy would have put it
// and its output rows are the rows of the whole job: results are bit-identical to the one-shot path.
// The offsets (needed by every chunk) are already on their way on the handle's stream; the caller holds
// the handle's lock. Host buffers should be pinned for the copies to overlap (pageable ones still work).
constexpr int64_t TOKENIZE_PIPE_CHUNK = 65536;

static int tokenize_pipelined(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n, int64_t bytes,
                              int32_t max_len, int32_t* ids, uint8_t* mask, int32_t* d_ids, uint8_t* d_mask,
                              b2_metrics* out) {
  constexpr int S = 3;
  for (int l = 0; l < S; ++l) {
    if (!h->copy_stream[l]) B2_CUDA_OK(cudaStreamCreateWithFlags(&h->copy_stream[l], cudaStreamNonBlocking));
    if (!h->pin_ev[l][0]) B2_CUDA_OK(cudaEventCreateWithFlags(&h->pin_ev[l][0], cudaEventDisableTiming));
  }
  cudaStream_t st = h->stream;
  B2_CUDA_OK(cudaEventRecord(h->ev_begin, st));
  B2_CUDA_OK(cudaEventRecord(h->ev_run, st));  // behind the upload of the offsets
  int rc = B2_OK;
  int64_t launches = 0;
  for (int64_t a = 0, c = 0; a < n; a += TOKENIZE_PIPE_CHUNK, ++c) {
    const int64_t cnt = std::min<int64_t>(TOKENIZE_PIPE_CHUNK, n - a);
    const int64_t lo = offsets[a], hi = offsets[a + cnt];
    if (lo < 0 || hi < lo || hi > bytes) {
      set_last_error("text input: offsets[%lld..%lld] are not non-decreasing within [0, %lld]",
                     static_cast<long long>(a), static_cast<long long>(a + cnt), static_cast<long long>(bytes));
      rc = B2_ERR_ARG;
      break;
    }
    cudaStream_t s = h->copy_stream[c % S];
    if (c < S) B2_CUDA_OK(cudaStreamWaitEvent(s, h->ev_run, 0));
    if (hi > lo) B2_CUDA_OK(cudaMemcpyAsync(h->st_text + lo, utf8 + lo, size_t(hi - lo), cudaMemcpyHostToDevice, s));
    if ((rc = tokenizer_run(s, h->tok, h->st_text, h->st_off + a, static_cast<int>(cnt), max_len,
                            d_ids + a * max_len, d_mask ? d_mask + a * max_len : nullptr, nullptr, h->num_sms)))
      break;
    ++launches;
    B2_CUDA_OK(cudaMemcpyAsync(ids + a * max_len, d_ids + a * max_len, size_t(cnt) * max_len * 4,
                               cudaMemcpyDeviceToHost, s));
    if (mask)
      B2_CUDA_OK(cudaMemcpyAsync(mask + a * max_len, d_mask + a * max_len, size_t(cnt) * max_len,
                                 cudaMemcpyDeviceToHost, s));
  }
  for (int l = 0; l < S; ++l) {   // the handle's stream waits for the side streams, also on the error path
    B2_CUDA_OK(cudaEventRecord(h->pin_ev[l][0], h->copy_stream[l]));
    B2_CUDA_OK(cudaStreamWaitEvent(st, h->pin_ev[l][0], 0));
  }
  B2_CUDA_OK(cudaEventRecord(h->ev_end, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  if (rc) return rc;
  B2_CUDA_OK(cudaGetLastError());
  float ms = 0.f;
  B2_CUDA_OK(cudaEventElapsedTime(&ms, h->ev_begin, h->ev_end));
  KindAcc acc;
  acc.launches = launches;
  fill_metrics(out, h, acc, n, 0, n * max_len, 1, 0, 0, ms);
  return B2_OK;
}

int b2_tokenize(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n, int32_t max_len,
                int32_t* ids, uint8_t* mask, b2_metrics* metrics) {
  if (!h) { set_last_error("b2_tokenize: NULL handle"); return B2_ERR_ARG; }
  int64_t bytes = 0;
  size_t out_ids = 0, out_mask = 0;
  bool pipelined = false;
  {
    std::lock_guard<std::mutex> lock(h->mu);
    DeviceGuard guard(h->device);
    if (max_len < 2) { set_last_error("tokenize: max_len must be >= 2"); return B2_ERR_ARG; }
    if (!h->has_tokenizer) {
      set_last_error("this handle was created without a tokenizer (b2_config.vocab_path was NULL)");
      return B2_ERR_ARG;
    }
    if (n > 0 && !ids) { set_last_error("tokenize: ids is NULL"); return B2_ERR_ARG; }
    int rc;
    pipelined = n >= 2 * TOKENIZE_PIPE_CHUNK && !h->profiling;
    if ((rc = stage_text(h, utf8, offsets, n, &bytes, !pipelined))) return rc;
    out_ids = size_t(n) * max_len * 4;
    out_mask = mask ? size_t(n) * max_len : 0;
    if ((rc = grow(&h->st_out, &h->st_out_cap, out_ids + out_mask + 64))) return rc;
  }
  int32_t* d_ids = static_cast<int32_t*>(h->st_out);
  uint8_t* d_mask = mask ? reinterpret_cast<uint8_t*>(h->st_out) + out_ids : nullptr;
  b2_metrics local;
  memset(&local, 0, sizeof(local));
  local.struct_size = sizeof(local);
  if (pipelined) {
    // large jobs: upload, kernel and download of successive chunks overlap on three streams
    std::lock_guard<std::mutex> lock(h->mu);
    DeviceGuard guard(h->device);
    int rc = tokenize_pipelined(h, utf8, offsets, n, bytes, max_len, ids, mask, d_ids, d_mask, &local);
    if (rc) return rc;
    local.h2d_bytes = bytes + (n + 1) * 8;
    local.d2h_bytes = out_ids + out_mask;
    if (metrics && metrics->struct_size > 0) {
      const int32_t keep = metrics->struct_size;
      memcpy(metrics, &local, std::min<size_t>(sizeof(local), keep));
      metrics->struct_size = keep;
    }
    return B2_OK;
  }
  int rc = b2_tokenize_device(h, h->st_text, h->st_off, n, bytes, max_len, d_ids, d_mask, &local);
  if (rc) return rc;
  {
    std::lock_guard<std::mutex> lock(h->mu);
    DeviceGuard guard(h->device);
    if (out_ids) B2_CUDA_OK(cudaMemcpy(ids, d_ids, out_ids, cudaMemcpyDeviceToHost));
    if (out_mask) B2_CUDA_OK(cudaMemcpy(mask, d_mask, out_mask, cudaMemcpyDeviceToHost));
  }
  local.h2d_bytes = bytes + (n + 1) * 8;
  local.d2h_bytes = out_ids + out_mask;
  if (metrics && metrics->struct_size > 0) {
    const int32_t keep = metrics->struct_size;
    memcpy(metrics, &local, std::min<size_t>(sizeof(local), keep));
    metrics->struct_size = keep;
  }
  return B2_OK;
}

int b2_classify_text_device(b2_handle* h, const uint8_t* d_utf8, const int64_t* d_offsets, int64_t n,
                            int64_t n_bytes, int32_t max_len, int32_t topk, int32_t* d_top_idx,
                            float* d_top_score, int32_t* d_labels, int64_t* d_histogram,
                            float* d_logits, b2_metrics* metrics) {
  (void)n_bytes;
  if (!h) { set_last_error("b2_classify_text_device: NULL handle"); return B2_ERR_ARG; }
  if (n > 0 && (!d_utf8 || !d_offsets)) {
    set_last_error("classify: NULL text / offsets");
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  return classify_device(h, d_utf8, d_offsets, nullptr, nullptr, n, max_len, topk, d_top_idx,
                         d_top_score, d_labels, d_histogram, d_logits, metrics, 0, 0);
}

// outputs staged on the device in one block: top_idx | top_score | labels | hist | logits
struct OutLayout {
  size_t top_idx, top_score, labels, hist, logits, total;
};
static OutLayout out_layout(int64_t n, int topk, int C, bool want_logits) {
  OutLayout L;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~size_t(255); return at; };
  L.top_idx = take(size_t(n) * topk * 4);
  L.top_score = take(size_t(n) * topk * 4);
  L.labels = take(size_t(n) * 4);
  L.hist = take(size_t(C) * 8);
  L.logits = take(want_logits ? size_t(n) * C * 4 : 0);
  L.total = o + 256;
  return L;
}

static int classify_to_host(b2_handle* h, const uint8_t* d_text, const int64_t* d_off, const int32_t* d_ids,
                            const uint8_t* d_mask, int64_t n, int32_t max_len, int32_t topk,
                            int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                            float* logits, b2_metrics* metrics, int64_t h2d);

// the caller holds the handle's lock and has made its device current
static int classify_host_locked(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, const int32_t* ids,
                                const uint8_t* mask, int64_t n, int32_t max_len, int32_t topk,
                                int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                                float* logits, b2_metrics* metrics) {
  int rc;
  if ((rc = check_len(h, max_len, true))) return rc;
  const int C = h->mh.num_labels;
  if (topk < 1 || topk > C) {
    set_last_error("topk %d is outside [1, num_labels = %d]", topk, C);
    return B2_ERR_ARG;
  }
  if (n < 0 || n > 0x7fffffff || (n > 0 && (!top_idx || !top_score))) {
    set_last_error("classify: n out of range or top_idx / top_score is NULL");
    return B2_ERR_ARG;
  }
  int64_t h2d = 0;
  const uint8_t* d_text = nullptr;
  const int64_t* d_off = nullptr;
  const int32_t* d_ids = nullptr;
  const uint8_t* d_mask = nullptr;
  if (utf8 || offsets) {
    int64_t bytes = 0;
    if ((rc = stage_text(h, utf8, offsets, n, &bytes))) return rc;
    d_text = h->st_text; d_off = h->st_off;
    h2d = bytes + (n + 1) * 8;
  } else {
    if (n > 0 && (!ids || !mask)) { set_last_error("classify: NULL ids / mask"); return B2_ERR_ARG; }
    const size_t ib = size_t(n) * max_len * 4, mb = size_t(n) * max_len;
    if ((rc = grow(reinterpret_cast<void**>(&h->st_text), &h->st_text_cap, ib + mb + 64))) return rc;
    if (ib) B2_CUDA_OK(cudaMemcpyAsync(h->st_text, ids, ib, cudaMemcpyHostToDevice, h->stream));
    if (mb) B2_CUDA_OK(cudaMemcpyAsync(h->st_text + ib, mask, mb, cudaMemcpyHostToDevice, h->stream));
    d_ids = reinterpret_cast<const int32_t*>(h->st_text);
    d_mask = h->st_text + ib;
    h2d = ib + mb;
  }
  return classify_to_host(h, d_text, d_off, d_ids, d_mask, n, max_len, topk, top_idx, top_score, labels,
                          histogram, logits, metrics, h2d);
}

// Run classify_device with the outputs staged in device memory, then copy them to the caller's host
// buffers. Inputs are device pointers; the caller holds the handle's lock and has checked topk.
static int classify_to_host(b2_handle* h, const uint8_t* d_text, const int64_t* d_off, const int32_t* d_ids,
                            const uint8_t* d_mask, int64_t n, int32_t max_len, int32_t topk,
                            int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                            float* logits, b2_metrics* metrics, int64_t h2d) {
  int rc;
  const int C = h->mh.num_labels;
  const OutLayout L = out_layout(n, topk, C, logits != nullptr);
  if ((rc = grow(&h->st_out, &h->st_out_cap, L.total))) return rc;
  uint8_t* ob = static_cast<uint8_t*>(h->st_out);
  int64_t d2h = size_t(n) * topk * 8 + (labels ? n * 4 : 0) + (histogram ? C * 8 : 0) +
                (logits ? size_t(n) * C * 4 : 0);
  rc = classify_device(h, d_text, d_off, d_ids, d_mask, n, max_len, topk,
                       reinterpret_cast<int32_t*>(ob + L.top_idx),
                       reinterpret_cast<float*>(ob + L.top_score),
                       reinterpret_cast<int32_t*>(ob + L.labels),
                       reinterpret_cast<int64_t*>(ob + L.hist),
                       logits ? reinterpret_cast<float*>(ob + L.logits) : nullptr, metrics, h2d, d2h);
  if (rc) return rc;
  if (n > 0) {
    B2_CUDA_OK(cudaMemcpy(top_idx, ob + L.top_idx, size_t(n) * topk * 4, cudaMemcpyDeviceToHost));
    B2_CUDA_OK(cudaMemcpy(top_score, ob + L.top_score, size_t(n) * topk * 4, cudaMemcpyDeviceToHost));
    if (labels) B2_CUDA_OK(cudaMemcpy(labels, ob + L.labels, size_t(n) * 4, cudaMemcpyDeviceToHost));
    if (logits) B2_CUDA_OK(cudaMemcpy(logits, ob + L.logits, size_t(n) * C * 4, cudaMemcpyDeviceToHost));
  }
  if (histogram) B2_CUDA_OK(cudaMemcpy(histogram, ob + L.hist, size_t(C) * 8, cudaMemcpyDeviceToHost));
  return B2_OK;
}

static int classify_host(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, const int32_t* ids,
                         const uint8_t* mask, int64_t n, int32_t max_len, int32_t topk,
                         int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                         float* logits, b2_metrics* metrics) {
  if (!h) { set_last_error("classify: NULL handle"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  return classify_host_locked(h, utf8, offsets, ids, mask, n, max_len, topk, top_idx, top_score, labels,
                              histogram, logits, metrics);
}

int b2_classify_text(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                     int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score,
                     int32_t* labels, int64_t* histogram, float* logits, b2_metrics* metrics) {
  if (!offsets) { set_last_error("b2_classify_text: offsets is NULL"); return B2_ERR_ARG; }
  static const uint8_t empty = 0;
  return classify_host(h, utf8 ? utf8 : &empty, offsets, nullptr, nullptr, n, max_len, topk, top_idx,
                       top_score, labels, histogram, logits, metrics);
}

int b2_classify_ids(b2_handle* h, const int32_t* ids, const uint8_t* mask, int64_t n, int32_t max_len,
                    int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                    int64_t* histogram, float* logits, b2_metrics* metrics) {
  return classify_host(h, nullptr, nullptr, ids, mask, n, max_len, topk, top_idx, top_score, labels,
                       histogram, logits, metrics);
}

int b2_debug_hidden_states(b2_handle* h, const int32_t* ids, const uint8_t* mask, int64_t n, int32_t max_len,
                           float* hidden, float* logits) {
  if (!h || !hidden) { set_last_error("b2_debug_hidden_states: NULL handle or output"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  int rc;
  if ((rc = check_len(h, max_len, true))) return rc;
  if (n < 1 || n > h->max_batch) {
    set_last_error("b2_debug_hidden_states: n = %lld is outside [1, max_batch = %d] (one batch only)",
                   static_cast<long long>(n), h->max_batch);
    return B2_ERR_LIMIT;
  }
  const size_t floats = size_t(h->mh.layers + 1) * size_t(n) * max_len * h->mh.hidden;
  float* d = nullptr;
  B2_CUDA_OK(cudaMalloc(&d, floats * sizeof(float)));
  struct Drop {  // whatever happens, the handle goes back to its normal state
    b2_handle* h; float* d;
    ~Drop() { h->dbg_hidden = nullptr; cudaFree(d); }
  } drop{h, d};
  B2_CUDA_OK(cudaMemsetAsync(d, 0, floats * sizeof(float), h->stream));
  h->dbg_hidden = d;
  std::vector<int32_t> top_idx(n);
  std::vector<float> top_score(n);
  if ((rc = classify_host_locked(h, nullptr, nullptr, ids, mask, n, max_len, 1, top_idx.data(), top_score.data(),
                                 nullptr, nullptr, logits, nullptr))) return rc;
  B2_CUDA_OK(cudaMemcpy(hidden, d, floats * sizeof(float), cudaMemcpyDeviceToHost));
  return B2_OK;
}

int b2_chunk_count(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                   int64_t* counts) {
  if (!h || !offsets || (n > 0 && !counts)) { set_last_error("b2_chunk_count: NULL argument"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  static const uint8_t empty = 0;
  int64_t bytes = 0;
  int rc;
  if ((rc = stage_text(h, utf8 ? utf8 : &empty, offsets, n, &bytes))) return rc;
  if ((rc = grow(&h->st_out, &h->st_out_cap, size_t(n) * 8 + 64))) return rc;
  if ((rc = chunk_count(h->stream, h->st_text, h->st_off, static_cast<int>(n),
                        static_cast<int64_t*>(h->st_out)))) return rc;
  B2_CUDA_OK(cudaStreamSynchronize(h->stream));
  if (n) B2_CUDA_OK(cudaMemcpy(counts, h->st_out, size_t(n) * 8, cudaMemcpyDeviceToHost));
  return B2_OK;
}

int b2_chunk_bounds(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                    int64_t chunk_size, const int64_t* chunk_base, int64_t n_chunks, int64_t* bounds) {
  if (!h || !offsets || (n > 0 && !chunk_base) || (n_chunks > 0 && !bounds) || chunk_size <= 0) {
    set_last_error("b2_chunk_bounds: NULL argument or chunk_size <= 0");
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  static const uint8_t empty = 0;
  int64_t bytes = 0;
  int rc;
  if ((rc = stage_text(h, utf8 ? utf8 : &empty, offsets, n, &bytes))) return rc;
  const size_t base_bytes = (size_t(n) * 8 + 255) & ~size_t(255);
  if ((rc = grow(&h->st_out, &h->st_out_cap, base_bytes + size_t(n_chunks) * 8 + 64))) return rc;
  int64_t* d_base = static_cast<int64_t*>(h->st_out);
  int64_t* d_bounds = reinterpret_cast<int64_t*>(static_cast<uint8_t*>(h->st_out) + base_bytes);
  if (n) B2_CUDA_OK(cudaMemcpyAsync(d_base, chunk_base, size_t(n) * 8, cudaMemcpyHostToDevice, h->stream));
  if ((rc = chunk_bounds(h->stream, h->st_text, h->st_off, static_cast<int>(n), chunk_size, d_base,
                         d_bounds))) return rc;
  B2_CUDA_OK(cudaStreamSynchronize(h->stream));
  if (n_chunks) B2_CUDA_OK(cudaMemcpy(bounds, d_bounds, size_t(n_chunks) * 8, cudaMemcpyDeviceToHost));
  return B2_OK;
}

}  // extern "C"

// ================================================================================================
// CSV: row index, field extraction and the csv -> tokenize -> classify chain
// ================================================================================================
// Counterpart of _read_csv_shard (ops/csv_shard.py:9-26). The reference re-parses the file from the
// top with csv.DictReader for every shard; here the file is indexed once (b2_csv_open) and shards
// are cut from the index.

struct b2_csv {
  b2_handle* h = nullptr;
  uint8_t* d_bytes = nullptr;
  int64_t n = 0, data_offset = 0, rows = 0;
  int64_t *row_start = nullptr, *row_end = nullptr;
  bool has_header = false;
  std::string header_bytes;         // unquoted header names back to back
  std::vector<int64_t> header_end;  // end offset of every name in header_bytes
  // result of the last sizing pass, reused by b2_csv_rows_read
  int64_t sz_start = -1, sz_rows = -1, sz_fields = 0, sz_bytes = 0;
  int32_t sz_col = -1;
  int64_t *d_nf = nullptr, *d_nb = nullptr, *d_fbase = nullptr, *d_bbase = nullptr;
  void* d_scan = nullptr;
  int64_t sz_cap = 0;
  // staging for extracted content
  uint8_t* d_out = nullptr; size_t d_out_cap = 0;
  int64_t* d_fend = nullptr; size_t d_fend_cap = 0;
};

namespace b2 {
namespace {

constexpr int64_t CSV_FIELD_LIMIT = 131072;  // csv.field_size_limit() default, in code points

// The header is the reader's first record whatever it holds (csv.DictReader.fieldnames), so it is
// cut off on the host before the device scan starts: a handful of bytes, walked once.
struct HeaderScan {
  bool present = false;    // false: the data ended before any record did (empty file)
  bool too_long = false;   // a name exceeds the field size limit
  int64_t long_offset = -1;
  int64_t data_offset = 0; // first byte after the header's line end ("\r\n" counted as one)
};

HeaderScan scan_header(const uint8_t* p, int64_t n, std::string* names, std::vector<int64_t>* ends) {
  HeaderScan r;
  enum { SR, SF, IF, IQ, QQ } s = SR;
  int64_t flen = 0;
  auto add = [&](uint8_t b, int64_t pos) {
    if ((b & 0xc0) != 0x80) {
      if (flen >= CSV_FIELD_LIMIT && !r.too_long) { r.too_long = true; r.long_offset = pos; }
      ++flen;
    }
    names->push_back(static_cast<char>(b));
  };
  auto save = [&]() { ends->push_back(static_cast<int64_t>(names->size())); flen = 0; };
  int64_t pos = 0;
  for (; pos < n; ++pos) {
    const uint8_t b = p[pos];
    const bool eol = b == '\n' || b == '\r';
    if (eol && s != IQ) {
      if (s != SR) save();  // SR: an empty first line gives the reader's empty record: no names
      r.present = true;
      r.data_offset = pos + 1;
      if (b == '\r') {
        if (pos + 1 < n && p[pos + 1] == '\n') r.data_offset = pos + 2;
      }
      return r;
    }
    switch (s) {
      case SR: case SF:
        if (b == '"') s = IQ;
        else if (b == ',') { save(); s = SF; }
        else { add(b, pos); s = IF; }
        break;
      case IF:
        if (b == ',') { save(); s = SF; }
        else add(b, pos);
        break;
      case IQ:
        if (b == '"') s = QQ;
        else add(b, pos);
        break;
      case QQ:
        if (b == '"') { add(b, pos); s = IQ; }
        else if (b == ',') { save(); s = SF; }
        else { add(b, pos); s = IF; }
        break;
    }
  }
  // data ended inside the header record
  if (s != SR) { save(); r.present = true; }
  r.data_offset = n;
  return r;
}

// Host-to-device copy of a large pageable buffer (an mmap of the file). cudaMemcpy from pageable
// memory runs at the pace of one CPU thread staging through the driver's bounce buffer (≈ 11 GB/s
// measured); here COPY_LANES threads each stage their own chunks through a pair of pinned buffers
// and issue their own asynchronous copies on a side stream, so the staging memcpys run in parallel
// with each other and with the DMA. The handle's stream is made to wait for all lanes.
constexpr size_t PIN_CHUNK = size_t(4) << 20;
constexpr int64_t PIPELINED_COPY_MIN = int64_t(32) << 20;

int copy_to_device_pipelined(b2_handle* h, uint8_t* d_dst, const uint8_t* src, int64_t n) {
  constexpr int L = b2_handle::COPY_LANES;
  for (int l = 0; l < L; ++l) {
    if (!h->copy_stream[l]) B2_CUDA_OK(cudaStreamCreateWithFlags(&h->copy_stream[l], cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      if (!h->pin_buf[l][b]) B2_CUDA_OK(cudaHostAlloc(reinterpret_cast<void**>(&h->pin_buf[l][b]), PIN_CHUNK, 0));
      if (!h->pin_ev[l][b]) B2_CUDA_OK(cudaEventCreateWithFlags(&h->pin_ev[l][b], cudaEventDisableTiming));
    }
  }
  const int64_t chunks = (n + int64_t(PIN_CHUNK) - 1) / int64_t(PIN_CHUNK);
  std::atomic<int> failed{0};
  auto lane = [&](int l) {
    if (cudaSetDevice(h->device) != cudaSuccess) { failed = 1; return; }
    int use[2] = {0, 0};
    for (int64_t ck = l; ck < chunks && !failed; ck += L) {
      const int b = static_cast<int>((ck / L) & 1);
      const int64_t off = ck * int64_t(PIN_CHUNK);
      const size_t len = static_cast<size_t>(std::min<int64_t>(int64_t(PIN_CHUNK), n - off));
      if (use[b] && cudaEventSynchronize(h->pin_ev[l][b]) != cudaSuccess) { failed = 1; return; }
      memcpy(h->pin_buf[l][b], src + off, len);
      if (cudaMemcpyAsync(d_dst + off, h->pin_buf[l][b], len, cudaMemcpyHostToDevice, h->copy_stream[l]) !=
              cudaSuccess ||
          cudaEventRecord(h->pin_ev[l][b], h->copy_stream[l]) != cudaSuccess) {
        failed = 1;
        return;
      }
      use[b] = 1;
    }
  };
  std::thread workers[L];
  for (int l = 0; l < L; ++l) workers[l] = std::thread(lane, l);
  for (int l = 0; l < L; ++l) workers[l].join();
  if (failed) {
    set_last_error("b2_csv_open: staged host-to-device copy failed: %s", cudaGetErrorString(cudaGetLastError()));
    return B2_ERR_CUDA;
  }
  for (int l = 0; l < L; ++l) {
    cudaEvent_t done = h->pin_ev[l][0];  // any event of the lane: re-record it at the end of the lane's stream
    B2_CUDA_OK(cudaEventRecord(done, h->copy_stream[l]));
    B2_CUDA_OK(cudaStreamWaitEvent(h->stream, done, 0));
  }
  return B2_OK;
}

int csv_size_range(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t col) {
  if (c->sz_start == start_row && c->sz_rows == n_rows && c->sz_col == col) return B2_OK;
  b2_handle* h = c->h;
  c->sz_start = -1;
  if (!c->d_nf || n_rows > c->sz_cap) {
    for (void* p : {static_cast<void*>(c->d_nf), static_cast<void*>(c->d_nb), static_cast<void*>(c->d_fbase),
                    static_cast<void*>(c->d_bbase), c->d_scan})
      if (p) cudaFree(p);
    c->d_nf = c->d_nb = c->d_fbase = c->d_bbase = nullptr;
    c->d_scan = nullptr;
    c->sz_cap = 0;
    const int64_t cap = n_rows + n_rows / 8 + 256;
    B2_CUDA_OK(cudaMalloc(&c->d_nf, size_t(cap) * 8));
    B2_CUDA_OK(cudaMalloc(&c->d_nb, size_t(cap) * 8));
    B2_CUDA_OK(cudaMalloc(&c->d_fbase, size_t(cap + 1) * 8));
    B2_CUDA_OK(cudaMalloc(&c->d_bbase, size_t(cap + 1) * 8));
    B2_CUDA_OK(cudaMalloc(&c->d_scan, scan_i64_scratch_bytes(cap)));
    c->sz_cap = cap;
  }
  int rc;
  cudaStream_t st = h->stream;
  if ((rc = csv_rows_size(st, c->d_bytes, c->row_start, c->row_end, start_row, n_rows, col, c->d_nf,
                          c->d_nb))) return rc;
  if ((rc = exclusive_scan_i64(st, c->d_nf, n_rows, c->d_fbase, c->d_scan))) return rc;
  if ((rc = exclusive_scan_i64(st, c->d_nb, n_rows, c->d_bbase, c->d_scan))) return rc;
  B2_CUDA_OK(cudaMemcpyAsync(&c->sz_fields, c->d_fbase + n_rows, 8, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaMemcpyAsync(&c->sz_bytes, c->d_bbase + n_rows, 8, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  c->sz_start = start_row;
  c->sz_rows = n_rows;
  c->sz_col = col;
  return B2_OK;
}

// content of rows [start_row, start_row + n_rows) into c->d_out / c->d_fend (device)
int csv_extract_range(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t col) {
  int rc;
  if ((rc = csv_size_range(c, start_row, n_rows, col))) return rc;
  if ((rc = grow(reinterpret_cast<void**>(&c->d_out), &c->d_out_cap, size_t(c->sz_bytes) + 64))) return rc;
  // one leading slot so that d_fend doubles as the offsets array of the tokenizer (offsets[0] = 0)
  if ((rc = grow(reinterpret_cast<void**>(&c->d_fend), &c->d_fend_cap, size_t(c->sz_fields + 2) * 8)))
    return rc;
  cudaStream_t st = c->h->stream;
  B2_CUDA_OK(cudaMemsetAsync(c->d_fend, 0, 8, st));
  return csv_rows_write(st, c->d_bytes, c->row_start, c->row_end, start_row, n_rows, col, c->d_fbase,
                        c->d_bbase, c->d_out, c->d_fend + 1);
}

int csv_check_range(b2_csv* c, int64_t start_row, int64_t n_rows, const char* who) {
  if (!c) { set_last_error("%s: NULL csv", who); return B2_ERR_ARG; }
  if (start_row < 0 || n_rows < 0 || start_row > c->rows || n_rows > c->rows - start_row) {
    set_last_error("%s: rows [%lld, %lld) are outside the file's %lld data rows", who,
                   static_cast<long long>(start_row), static_cast<long long>(start_row + n_rows),
                   static_cast<long long>(c->rows));
    return B2_ERR_ARG;
  }
  return B2_OK;
}

}  // namespace
}  // namespace b2

extern "C" {

void b2_csv_close(b2_csv* c) {
  if (!c) return;
  {
    std::lock_guard<std::mutex> lock(c->h->mu);
    DeviceGuard guard(c->h->device);
    for (void* p : {static_cast<void*>(c->d_bytes), static_cast<void*>(c->row_start),
                    static_cast<void*>(c->row_end), static_cast<void*>(c->d_nf), static_cast<void*>(c->d_nb),
                    static_cast<void*>(c->d_fbase), static_cast<void*>(c->d_bbase), c->d_scan,
                    static_cast<void*>(c->d_out), static_cast<void*>(c->d_fend)})
      if (p) cudaFree(p);
  }
  delete c;
}

int b2_csv_open(b2_handle* h, const uint8_t* bytes, int64_t n_bytes, b2_csv** out, b2_csv_info* info) {
  if (!h || !out || n_bytes < 0 || (n_bytes > 0 && !bytes)) {
    set_last_error("b2_csv_open: NULL handle / bytes / out, or negative size");
    return B2_ERR_ARG;
  }
  *out = nullptr;
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  cudaStream_t st = h->stream;
  b2_csv* c = new b2_csv();
  c->h = h;
  c->n = n_bytes;
  struct Fail {  // frees the half-built object unless released
    b2_csv* c;
    ~Fail() {
      if (!c) return;
      for (void* p : {static_cast<void*>(c->d_bytes), static_cast<void*>(c->row_start),
                      static_cast<void*>(c->row_end)})
        if (p) cudaFree(p);
      delete c;
    }
  } fail{c};
  void* scratch = nullptr;
  struct FreeLater { void** p; ~FreeLater() { if (*p) cudaFree(*p); } } free_scratch{&scratch};

  B2_CUDA_OK(cudaMalloc(&c->d_bytes, size_t(n_bytes) + 64));
  B2_CUDA_OK(cudaEventRecord(h->ev_begin, st));
  if (n_bytes >= PIPELINED_COPY_MIN) {
    int rc_copy;
    if ((rc_copy = copy_to_device_pipelined(h, c->d_bytes, bytes, n_bytes))) return rc_copy;
  } else if (n_bytes) {
    B2_CUDA_OK(cudaMemcpyAsync(c->d_bytes, bytes, n_bytes, cudaMemcpyHostToDevice, st));
  }
  B2_CUDA_OK(cudaMemsetAsync(c->d_bytes + n_bytes, 0, 64, st));

  const HeaderScan hs = scan_header(bytes, n_bytes, &c->header_bytes, &c->header_end);
  c->has_header = hs.present;
  c->data_offset = hs.data_offset;

  B2_CUDA_OK(cudaMallocAsync(&scratch, csv_index_scratch_bytes(n_bytes) + 64, st));  // stream-ordered: no device-wide stall
  cudaEvent_t ev_copied = nullptr;
  B2_CUDA_OK(cudaEventCreate(&ev_copied));
  struct DropEvent { cudaEvent_t e; ~DropEvent() { cudaEventDestroy(e); } } drop_event{ev_copied};
  B2_CUDA_OK(cudaEventRecord(ev_copied, st));
  // results: [0] rows (i64) | [1] final state (i32) | [2] first invalid utf-8 | [3..4] field limit
  int64_t* d_res = static_cast<int64_t*>(scratch);
  void* d_index = static_cast<uint8_t*>(scratch) + 64;
  int rc;
  if ((rc = csv_count_rows(st, c->d_bytes, c->data_offset, n_bytes, d_index, d_res,
                           reinterpret_cast<int32_t*>(d_res + 1),
                           reinterpret_cast<unsigned long long*>(d_res + 2)))) return rc;
  int64_t res[5] = {0, 0, -1, -1, -1};
  B2_CUDA_OK(cudaMemcpyAsync(res, d_res, 24, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  c->rows = res[0];
  const int final_state = static_cast<int>(res[1] & 0xffffffff);
  B2_CUDA_OK(cudaMallocAsync(&c->row_start, size_t(c->rows + 1) * 8, st));
  B2_CUDA_OK(cudaMallocAsync(&c->row_end, size_t(c->rows + 1) * 8, st));
  if ((rc = csv_emit_rows(st, c->data_offset, n_bytes, d_index, c->row_start, c->row_end)))
    return rc;
  if (c->rows > 0 && final_state != 0)  // last row has no line end
    B2_CUDA_OK(cudaMemcpyAsync(c->row_end + c->rows - 1, &c->n, 8, cudaMemcpyHostToDevice, st));
  if ((rc = csv_field_limit(st, c->d_bytes, c->row_start, c->row_end, c->rows, CSV_FIELD_LIMIT,
                            reinterpret_cast<unsigned long long*>(d_res + 3)))) return rc;
  B2_CUDA_OK(cudaMemcpyAsync(res + 3, d_res + 3, 16, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaEventRecord(h->ev_end, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  float ms = 0.f, index_ms = 0.f;
  B2_CUDA_OK(cudaEventElapsedTime(&ms, h->ev_begin, h->ev_end));
  B2_CUDA_OK(cudaEventElapsedTime(&index_ms, ev_copied, h->ev_end));
  if (info) {
    const int32_t want = info->struct_size;
    b2_csv_info v{};
    v.struct_size = sizeof(b2_csv_info);
    v.has_header = hs.present ? 1 : 0;
    v.header_fields = static_cast<int32_t>(c->header_end.size());
    v.header_bytes = static_cast<int64_t>(c->header_bytes.size());
    v.n_bytes = n_bytes;
    v.data_offset = c->data_offset;
    v.rows = c->rows;
    v.first_invalid_utf8 = res[2];  // 2^64-1 reads as -1
    v.header_field_too_long = hs.too_long ? 1 : 0;
    v.header_long_offset = hs.long_offset;
    v.long_field_row = res[3];
    v.long_field_offset = res[4];
    v.end_state = final_state;
    v.gpu_ms = ms;
    v.index_ms = index_ms;
    v.kernel_launches = 6;
    memcpy(info, &v, std::min<size_t>(sizeof(v), want > 0 ? size_t(want) : sizeof(v)));
  }
  fail.c = nullptr;
  *out = c;
  return B2_OK;
}

int b2_csv_header(b2_csv* c, uint8_t* names, int64_t names_cap, int64_t* name_end, int32_t max_names) {
  if (!c) { set_last_error("b2_csv_header: NULL csv"); return B2_ERR_ARG; }
  if (names_cap < static_cast<int64_t>(c->header_bytes.size()) ||
      max_names < static_cast<int32_t>(c->header_end.size())) {
    set_last_error("b2_csv_header: buffers too small (need %zu bytes, %zu names)", c->header_bytes.size(),
                   c->header_end.size());
    return B2_ERR_LIMIT;
  }
  if (!c->header_bytes.empty()) memcpy(names, c->header_bytes.data(), c->header_bytes.size());
  for (size_t i = 0; i < c->header_end.size(); ++i) name_end[i] = c->header_end[i];
  return B2_OK;
}

int b2_csv_row_spans(b2_csv* c, int64_t start_row, int64_t n_rows, int64_t* row_start, int64_t* row_end) {
  int rc;
  if ((rc = csv_check_range(c, start_row, n_rows, "b2_csv_row_spans"))) return rc;
  if (n_rows == 0) return B2_OK;
  std::lock_guard<std::mutex> lock(c->h->mu);
  DeviceGuard guard(c->h->device);
  if (row_start)
    B2_CUDA_OK(cudaMemcpy(row_start, c->row_start + start_row, size_t(n_rows) * 8, cudaMemcpyDeviceToHost));
  if (row_end)
    B2_CUDA_OK(cudaMemcpy(row_end, c->row_end + start_row, size_t(n_rows) * 8, cudaMemcpyDeviceToHost));
  return B2_OK;
}

int b2_csv_rows_size(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t column, int64_t* n_fields,
                     int64_t* n_bytes) {
  int rc;
  if ((rc = csv_check_range(c, start_row, n_rows, "b2_csv_rows_size"))) return rc;
  std::lock_guard<std::mutex> lock(c->h->mu);
  DeviceGuard guard(c->h->device);
  if ((rc = csv_size_range(c, start_row, n_rows, column))) return rc;
  if (n_fields) *n_fields = c->sz_fields;
  if (n_bytes) *n_bytes = c->sz_bytes;
  return B2_OK;
}

int b2_csv_rows_read(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t column, uint8_t* bytes,
                     int64_t* field_end, int64_t* row_field_end) {
  int rc;
  if ((rc = csv_check_range(c, start_row, n_rows, "b2_csv_rows_read"))) return rc;
  std::lock_guard<std::mutex> lock(c->h->mu);
  DeviceGuard guard(c->h->device);
  if ((rc = csv_extract_range(c, start_row, n_rows, column))) return rc;
  cudaStream_t st = c->h->stream;
  if (bytes && c->sz_bytes)
    B2_CUDA_OK(cudaMemcpyAsync(bytes, c->d_out, size_t(c->sz_bytes), cudaMemcpyDeviceToHost, st));
  if (field_end && c->sz_fields)
    B2_CUDA_OK(cudaMemcpyAsync(field_end, c->d_fend + 1, size_t(c->sz_fields) * 8, cudaMemcpyDeviceToHost, st));
  if (row_field_end && n_rows)
    B2_CUDA_OK(cudaMemcpyAsync(row_field_end, c->d_fbase + 1, size_t(n_rows) * 8, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  return B2_OK;
}

int b2_csv_classify(b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows, int32_t max_len,
                    int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                    float* logits, b2_metrics* metrics) {
  int rc;
  if ((rc = csv_check_range(c, start_row, n_rows, "b2_csv_classify"))) return rc;
  if (column < 0) { set_last_error("b2_csv_classify: column must be >= 0"); return B2_ERR_ARG; }
  b2_handle* h = c->h;
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  if ((rc = check_len(h, max_len, true))) return rc;
  if (!h->has_tokenizer) {
    set_last_error("b2_csv_classify: the handle has no tokenizer");
    return B2_ERR_ARG;
  }
  const int C = h->mh.num_labels;
  if (topk < 1 || topk > C) {
    set_last_error("topk %d is outside [1, num_labels = %d]", topk, C);
    return B2_ERR_ARG;
  }
  if (n_rows > 0x7fffffff || (n_rows > 0 && (!top_idx || !top_score))) {
    set_last_error("b2_csv_classify: too many rows or top_idx / top_score is NULL");
    return B2_ERR_ARG;
  }
  if ((rc = csv_extract_range(c, start_row, n_rows, column))) return rc;
  return classify_to_host(h, c->d_out, c->d_fend, nullptr, nullptr, n_rows, max_len, topk, top_idx,
                          top_score, labels, histogram, logits, metrics, 0);
}

}  // extern "C"

// ---- hooks for the multi-device group (b2_group.cu, b2_internal.h) -------------------------------------
#include "b2_internal.h"
namespace b2 {

int handle_device(const b2_handle* h) { return h->device; }
int handle_num_labels(const b2_handle* h) { return h->has_model ? static_cast<int>(h->mh.num_labels) : 0; }
bool handle_has_model(const b2_handle* h) { return h->has_model; }

int classify_resident_text(b2_handle* h, const uint8_t* d_text, const int64_t* d_off, int64_t n,
                           int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score,
                           int32_t* labels, int64_t* histogram, float* logits, b2_metrics* metrics) {
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  int rc;
  if ((rc = check_len(h, max_len, true))) return rc;
  const int C = h->mh.num_labels;
  if (topk < 1 || topk > C) {
    set_last_error("topk %d is outside [1, num_labels = %d]", topk, C);
    return B2_ERR_ARG;
  }
  if (n < 0 || n > 0x7fffffff || (n > 0 && (!top_idx || !top_score || !d_off || !d_text))) {
    set_last_error("classify: n out of range or NULL text / offsets / top_idx / top_score");
    return B2_ERR_ARG;
  }
  // a non-null text pointer is what selects the tokenizer path; an empty job never reads through it
  const uint8_t* text = d_text ? d_text : reinterpret_cast<const uint8_t*>(h->d_tokens);
  return classify_to_host(h, text, d_off, nullptr, nullptr, n, max_len, topk, top_idx, top_score, labels,
                          histogram, logits, metrics, 0);
}

b2_handle* csv_owner(b2_csv* c) { return c ? c->h : nullptr; }

int csv_column_extract(b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows, int64_t* n_bytes) {
  int rc;
  if ((rc = csv_check_range(c, start_row, n_rows, "b2_group_csv_classify"))) return rc;
  if (column < 0) { set_last_error("b2_group_csv_classify: column must be >= 0"); return B2_ERR_ARG; }
  std::lock_guard<std::mutex> lock(c->h->mu);
  DeviceGuard guard(c->h->device);
  if ((rc = csv_extract_range(c, start_row, n_rows, column))) return rc;
  *n_bytes = c->sz_bytes;
  return B2_OK;
}

int csv_column_copy_to(b2_csv* c, int dst_device, uint8_t* d_text, int64_t* d_off, int64_t n_rows,
                       int64_t n_bytes) {
  std::lock_guard<std::mutex> lock(c->h->mu);
  DeviceGuard guard(c->h->device);
  cudaStream_t st = c->h->stream;
  if (n_bytes)
    B2_CUDA_OK(cudaMemcpyPeerAsync(d_text, dst_device, c->d_out, c->h->device, size_t(n_bytes), st));
  B2_CUDA_OK(cudaMemcpyPeerAsync(d_off, dst_device, c->d_fend, c->h->device, size_t(n_rows + 1) * 8, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  return B2_OK;
}

}  // namespace b2

// Developer aid (tests/tools): a block of mapped host memory in which the first bounded mbarrier wait
// that gives up on the current device records itself before it traps — [0] 0xB2B2B2B2 when set,
// [1] blockDim.x << 32 | blockIdx.x, [2] barrier smem address << 32 | threadIdx.x,
// [3] gridDim.x << 32 | parity, [4] %globaltimer. Host memory survives the trapped context.
extern "C" unsigned long long* b2_debug_trap_log(void) {
  static unsigned long long* host = nullptr;
  if (!host) {
    if (cudaHostAlloc(reinterpret_cast<void**>(&host), 64, cudaHostAllocMapped) != cudaSuccess) return nullptr;
    memset(host, 0, 64);
  }
  unsigned long long* dev = nullptr;
  if (cudaHostGetDevicePointer(reinterpret_cast<void**>(&dev), host, 0) != cudaSuccess) return nullptr;
  if (b2::gemm_set_trap_log(dev) || b2::gemm2_set_trap_log(dev) || b2::attention_set_trap_log(dev)) return nullptr;
  return host;
}

// ================================================================================================
// b2_tokenize_noalloc — tokenize without allocating device memory
// ================================================================================================
// b2_tokenize stages the caller's host text on the device in buffers it grows with cudaMalloc when a
// job is larger than the last one. Some callers must not trigger that: a worker whose memory budget
// was handed out at start-up and who may not take more from the driver later, or a latency-bound
// loop that cannot afford the device-wide synchronisation an allocation or free implies. For them
// the staging area below is part of the library image instead: a fixed block of device memory that
// exists on every GPU from the moment the library is loaded there until the process exits, whether
// or not this entry point is ever called.
//
// Rules that follow from sharing one fixed block:
//   * a job whose text, offsets, ids and mask do not fit in B2_NOALLOC_SCRATCH_BYTES is refused;
//   * the block keeps whatever the previous call left in it, so the output region is cleared
//     before the kernel runs;
//   * two calls that overlap in time on the same GPU (different handles, different threads) would
//     write the same bytes — callers must serialise them;
//   * the first call on each GPU looks up the block's device address (a one-off set-up step);
//     make that first call outside any stream capture.
static __device__ __align__(256) uint8_t g_noalloc_scratch[B2_NOALLOC_SCRATCH_BYTES];

namespace {

std::mutex g_noalloc_mu;
uint8_t* g_noalloc_ptr[64] = {};  // per device: address of g_noalloc_scratch on that device

// `g_noalloc_scratch` in host code is only a handle for the runtime, not an address a kernel can
// use; the device address has to be asked for, once per device, with that device current.
int noalloc_scratch_for(int device, uint8_t** out) {
  if (device < 0 || device >= 64) {
    set_last_error("b2_tokenize_noalloc: device ordinal %d is out of range", device);
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // caller made `device` current
    if (e != cudaSuccess || !p) {
      set_last_error("b2_tokenize_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                     cudaGetErrorString(e));
      return B2_ERR_CUDA;
    }
    g_noalloc_ptr[device] = static_cast<uint8_t*>(p);
  }
  *out = g_noalloc_ptr[device];
  return B2_OK;
}

inline size_t up256(size_t v) { return (v + 255) & ~size_t(255); }

}  // namespace

extern "C" int b2_tokenize_noalloc(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                                   int32_t max_len, int32_t* ids, uint8_t* mask) {
  if (!h || !offsets || (n > 0 && !ids)) {
    set_last_error("b2_tokenize_noalloc: NULL handle, offsets or ids");
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(h->mu);
  DeviceGuard guard(h->device);
  if (!h->has_tokenizer) {
    set_last_error("this handle was created without a tokenizer (b2_config.vocab_path was NULL)");
    return B2_ERR_ARG;
  }
  if (max_len < 2 || n < 0 || n > 0x7fffffff) {
    set_last_error("b2_tokenize_noalloc: need max_len >= 2 and 0 <= n < 2^31");
    return B2_ERR_ARG;
  }
  const int64_t bytes = offsets[n];
  if (offsets[0] != 0 || bytes < 0 || (bytes > 0 && !utf8)) {
    set_last_error("b2_tokenize_noalloc: offsets must start at 0 and utf8 must not be NULL");
    return B2_ERR_ARG;
  }
  // Every part is held to the block size on its own before the parts are added up, so that no
  // product or sum of caller-supplied sizes can wrap around and slip under the limit.
  constexpr int64_t LIMIT = B2_NOALLOC_SCRATCH_BYTES;
  if (bytes > LIMIT || n > LIMIT / 8 || max_len > LIMIT / 4 || (n > 0 && n > LIMIT / 5 / max_len)) {
    set_last_error("b2_tokenize_noalloc: job (%lld text bytes, %lld items x max_len %d) cannot fit the fixed "
                   "block of %d bytes; split the job or use b2_tokenize",
                   (long long)bytes, (long long)n, max_len, B2_NOALLOC_SCRATCH_BYTES);
    return B2_ERR_LIMIT;
  }
  // The kernel reads item i at bytes [offsets[i], offsets[i+1]) of the staged text: a descending or
  // overshooting entry would send it outside the text part of the block (n is bounded by the check
  // above, so this pass is short).
  for (int64_t i = 0; i < n; ++i) {
    if (offsets[i] > offsets[i + 1] || offsets[i + 1] > bytes) {
      set_last_error("b2_tokenize_noalloc: offsets[%lld..%lld] are not non-decreasing within [0, %lld]",
                     (long long)i, (long long)i + 1, (long long)bytes);
      return B2_ERR_ARG;
    }
  }
  // layout of the fixed block: text | offsets | ids | mask, each part 256-byte aligned
  const size_t off_text = 0;
  const size_t off_offs = off_text + up256(static_cast<size_t>(bytes) + 16);
  const size_t off_ids = off_offs + up256(static_cast<size_t>(n + 1) * 8);
  const size_t off_mask = off_ids + up256(static_cast<size_t>(n) * max_len * 4);
  const size_t total = off_mask + up256(static_cast<size_t>(n) * max_len);
  if (total > B2_NOALLOC_SCRATCH_BYTES) {
    set_last_error("b2_tokenize_noalloc: job needs %zu bytes of staging (%lld text bytes, %lld items x "
                   "max_len %d) but the fixed block holds %d; split the job or use b2_tokenize",
                   total, (long long)bytes, (long long)n, max_len, B2_NOALLOC_SCRATCH_BYTES);
    return B2_ERR_LIMIT;
  }
  uint8_t* base = nullptr;
  int rc = noalloc_scratch_for(h->device, &base);
  if (rc) return rc;
  if (n == 0) return B2_OK;
  uint8_t* d_text = base + off_text;
  int64_t* d_offs = reinterpret_cast<int64_t*>(base + off_offs);
  int32_t* d_ids = reinterpret_cast<int32_t*>(base + off_ids);
  uint8_t* d_mask = base + off_mask;
  cudaStream_t st = h->stream;
  if (bytes) B2_CUDA_OK(cudaMemcpyAsync(d_text, utf8, bytes, cudaMemcpyHostToDevice, st));
  B2_CUDA_OK(cudaMemcpyAsync(d_offs, offsets, static_cast<size_t>(n + 1) * 8, cudaMemcpyHostToDevice, st));
  // the block still holds the previous call's rows: start the output region from zero
  B2_CUDA_OK(cudaMemsetAsync(base + off_ids, 0, total - off_ids, st));
  if ((rc = tokenizer_run(st, h->tok, d_text, d_offs, static_cast<int>(n), max_len, d_ids, d_mask, nullptr,
                          h->num_sms))) return rc;
  B2_CUDA_OK(cudaMemcpyAsync(ids, d_ids, static_cast<size_t>(n) * max_len * 4, cudaMemcpyDeviceToHost, st));
  if (mask) B2_CUDA_OK(cudaMemcpyAsync(mask, d_mask, static_cast<size_t>(n) * max_len, cudaMemcpyDeviceToHost, st));
  B2_CUDA_OK(cudaStreamSynchronize(st));
  B2_CUDA_OK(cudaGetLastError());
  return B2_OK;
}
