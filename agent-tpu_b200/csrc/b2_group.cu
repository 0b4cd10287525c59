// b2_group.cu — one job across several GPUs of one box from ONE process (include/b2_classify.h:
// b2_group_*).
//
// The reference's agent runs one task at a time, inline, in one process (app.py:31 MAX_TASKS=1,
// app.py:191 takes tasks[0], app.py:286 "run inline (no fork / no process pool)"), so the call it
// makes — tpu_ops.map_classify_tpu(payload) — has to fan a job out over the box's GPUs by itself. A
// group owns one b2_handle per device. A job's items are cut into contiguous slices, one per device
// (items are independent; no data-path collective); one host thread per device stages its slice,
// runs the same per-device pipeline as b2_classify_text and writes its results straight into the
// caller's arrays at the slice's position. The one exchange per job is an NCCL all-reduce (sum) of
// int64[num_labels + 8] — label histogram, additive counters, failure count — plus a max-reduce of
// the device times, over communicators made once with ncclCommInitAll.
//
// NCCL is looked up at run time (dlopen), not linked: a single-GPU deployment does not need it, and
// a process that has already loaded a libnccl.so.2 (PyTorch ships one) must keep using that copy.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "b2_internal.h"

namespace {

struct NcclApi {
  void* lib = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
};

std::mutex g_nccl_mu;
NcclApi g_nccl;

// Search order: a copy this process has already loaded (RTLD_NOLOAD), $B2_NCCL_PATH, the dynamic
// linker's default search.
int load_nccl(const NcclApi** out) {
  std::lock_guard<std::mutex> lock(g_nccl_mu);
  if (!g_nccl.lib) {
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    const char* env = getenv("B2_NCCL_PATH");
    if (!lib && env && env[0]) lib = dlopen(env, RTLD_NOW | RTLD_LOCAL);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!lib) {
      b2::set_last_error("b2_group_create: more than one device needs NCCL, and libnccl.so.2 cannot be loaded "
                         "(%s); set B2_NCCL_PATH", dlerror());
      return B2_ERR_IO;
    }
    NcclApi a;
    a.lib = lib;
    a.CommInitAll = reinterpret_cast<decltype(a.CommInitAll)>(dlsym(lib, "ncclCommInitAll"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(lib, "ncclAllReduce"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    a.GetVersion = reinterpret_cast<decltype(a.GetVersion)>(dlsym(lib, "ncclGetVersion"));
    if (!a.CommInitAll || !a.CommDestroy || !a.AllReduce || !a.GetErrorString) {
      b2::set_last_error("b2_group_create: libnccl.so.2 lacks ncclCommInitAll / ncclAllReduce");
      return B2_ERR_IO;
    }
    g_nccl = a;
  }
  *out = &g_nccl;
  return B2_OK;
}

constexpr int N_COUNTERS = 7;  // items, tokens, padded_tokens, batches, kernel_launches, h2d, d2h
constexpr int N_EXTRA = N_COUNTERS + 1;  // + failure count

}  // namespace

struct b2_group {
  std::vector<int> devices;
  std::vector<b2_handle*> members;
  const NcclApi* nccl = nullptr;
  std::vector<ncclComm_t> comms;
  int nccl_version = 0;
  int num_labels = 0;
  // per member, on its device: the vector that is all-reduced and the device time
  std::vector<int64_t*> d_vec;
  std::vector<double*> d_ms;
  // per member: staging for text that reaches it from another GPU (CSV chain)
  std::vector<uint8_t*> d_text;
  std::vector<size_t> d_text_cap;
  std::vector<int64_t*> d_off;
  std::vector<size_t> d_off_cap;
  std::mutex mu;  // one job at a time (the agent runs tasks inline: app.py:286)
  int64_t allreduces = 0;
};

namespace {

// The caller's current device is put back when a group call returns: member 0's slice runs on the
// calling thread and selects its own device.
struct CallerDevice {
  int prev = -1;
  CallerDevice() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
  ~CallerDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

struct SliceResult {
  int rc = B2_OK;
  std::string error;
  b2_metrics m;
  std::vector<int64_t> hist;
};

void fail(SliceResult* r, int rc) {
  r->rc = rc;
  r->error = b2::last_error();
}

// The job's one exchange: every member contributes (histogram | counters | failed?) and its device
// time; after it every member holds the totals, and member 0's copy is handed back. Always entered
// by every member thread, failed or not, so no thread is left waiting in the collective.
void reduce_member(b2_group* g, int i, SliceResult* r, std::vector<int64_t>* total, double* max_ms) {
  const int C = g->num_labels, count = C + N_EXTRA;
  std::vector<int64_t> vec(count, 0);
  if (r->rc == B2_OK) {
    for (int c = 0; c < C; ++c) vec[c] = r->hist[c];
    const int64_t counters[N_COUNTERS] = {r->m.items, r->m.tokens, r->m.padded_tokens, r->m.batches,
                                          r->m.kernel_launches, r->m.h2d_bytes, r->m.d2h_bytes};
    for (int k = 0; k < N_COUNTERS; ++k) vec[C + k] = counters[k];
  } else {
    vec[C + N_COUNTERS] = 1;
  }
  double ms = r->rc == B2_OK ? r->m.gpu_ms_total : 0.0;
  const int W = static_cast<int>(g->members.size());
  if (W == 1) {
    *total = vec;
    *max_ms = ms;
    return;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(static_cast<uintptr_t>(b2_get_stream(g->members[i])));
  cudaError_t e = cudaMemcpyAsync(g->d_vec[i], vec.data(), size_t(count) * 8, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(g->d_ms[i], &ms, 8, cudaMemcpyHostToDevice, st);
  ncclResult_t n1 = g->nccl->AllReduce(g->d_vec[i], g->d_vec[i], count, ncclInt64, ncclSum, g->comms[i], st);
  ncclResult_t n2 = g->nccl->AllReduce(g->d_ms[i], g->d_ms[i], 1, ncclFloat64, ncclMax, g->comms[i], st);
  if (e == cudaSuccess && i == 0) {
    total->assign(count, 0);
    e = cudaMemcpyAsync(total->data(), g->d_vec[i], size_t(count) * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(max_ms, g->d_ms[i], 8, cudaMemcpyDeviceToHost, st);
  }
  const cudaError_t es = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = es;
  if (r->rc == B2_OK && (e != cudaSuccess || n1 != ncclSuccess || n2 != ncclSuccess)) {
    b2::set_last_error("b2_group: histogram all-reduce failed on device %d: %s / %s", g->devices[i],
                       cudaGetErrorString(e), g->nccl->GetErrorString(n1 != ncclSuccess ? n1 : n2));
    fail(r, B2_ERR_CUDA);
  }
}

void add_kernel_times(b2_metrics* t, const b2_metrics& m) {
  t->ms_tokenize += m.ms_tokenize; t->ms_embed += m.ms_embed; t->ms_gemm_qkv += m.ms_gemm_qkv;
  t->ms_attention += m.ms_attention; t->ms_gemm_attn_out += m.ms_gemm_attn_out; t->ms_layernorm += m.ms_layernorm;
  t->ms_gemm_ffn1 += m.ms_gemm_ffn1; t->ms_gemm_ffn2 += m.ms_gemm_ffn2; t->ms_pooler += m.ms_pooler;
  t->ms_head += m.ms_head;
  t->n_tokenize += m.n_tokenize; t->n_embed += m.n_embed; t->n_gemm_qkv += m.n_gemm_qkv;
  t->n_attention += m.n_attention; t->n_gemm_attn_out += m.n_gemm_attn_out; t->n_layernorm += m.n_layernorm;
  t->n_gemm_ffn1 += m.n_gemm_ffn1; t->n_gemm_ffn2 += m.n_gemm_ffn2; t->n_pooler += m.n_pooler;
  t->n_head += m.n_head;
}

void put_metrics(b2_metrics* out, const b2_metrics& v) {
  if (!out || out->struct_size <= 0) return;
  const int32_t keep = out->struct_size;
  memcpy(out, &v, std::min<size_t>(sizeof(v), static_cast<size_t>(keep)));
  out->struct_size = keep;
}

// Collect what the member threads left: first error wins; totals from the all-reduce.
int finish_job(b2_group* g, std::vector<SliceResult>& res, const std::vector<int64_t>& total, double max_ms,
               int64_t* histogram, b2_metrics* metrics, b2_metrics* per_device) {
  const int W = static_cast<int>(g->members.size()), C = g->num_labels;
  for (int i = 0; i < W; ++i) {
    if (res[i].rc != B2_OK) {
      b2::set_last_error("device %d: %s", g->devices[i], res[i].error.c_str());
      return res[i].rc;
    }
  }
  if (static_cast<int>(total.size()) != C + N_EXTRA || total[C + N_COUNTERS] != 0) {
    b2::set_last_error("b2_group: the job's all-reduce reports a failed slice");
    return B2_ERR_CUDA;
  }
  if (histogram) for (int c = 0; c < C; ++c) histogram[c] = total[c];
  b2_metrics t;
  memset(&t, 0, sizeof(t));
  t.struct_size = sizeof(t);
  t.items = total[C + 0]; t.tokens = total[C + 1]; t.padded_tokens = total[C + 2]; t.batches = total[C + 3];
  t.kernel_launches = total[C + 4]; t.h2d_bytes = total[C + 5]; t.d2h_bytes = total[C + 6];
  t.gpu_ms_total = max_ms;
  for (int i = 0; i < W; ++i) {
    t.profiled |= res[i].m.profiled;
    add_kernel_times(&t, res[i].m);
    if (per_device) put_metrics(per_device + i, res[i].m);
  }
  put_metrics(metrics, t);
  if (W > 1) g->allreduces += 2;
  return B2_OK;
}

int grow_dev(void** p, size_t* cap, size_t need) {
  if (need <= *cap) return B2_OK;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  const size_t want = need + need / 4 + 4096;
  B2_CUDA_OK(cudaMalloc(p, want));
  *cap = want;
  return B2_OK;
}

int check_outputs(const b2_group* g, int64_t n, int32_t topk, const int32_t* top_idx, const float* top_score) {
  if (!g) { b2::set_last_error("b2_group: NULL group"); return B2_ERR_ARG; }
  if (n < 0 || n > 0x7fffffff || (n > 0 && (!top_idx || !top_score))) {
    b2::set_last_error("b2_group: n out of range or top_idx / top_score is NULL");
    return B2_ERR_ARG;
  }
  if (topk < 1 || topk > g->num_labels) {
    b2::set_last_error("topk %d is outside [1, num_labels = %d]", topk, g->num_labels);
    return B2_ERR_ARG;
  }
  return B2_OK;
}

}  // namespace

extern "C" {

int b2_group_create(const b2_config* cfg, const int32_t* devices, int32_t n_devices, b2_group** out) {
  if (!cfg || !out || cfg->struct_size < static_cast<int32_t>(sizeof(b2_config))) {
    b2::set_last_error("b2_group_create: config is NULL or has the wrong struct_size");
    return B2_ERR_ARG;
  }
  *out = nullptr;
  const int ndev = b2_device_count();
  if (ndev <= 0) {
    b2::set_last_error("No CUDA device detected (cudaGetDeviceCount returned 0).");
    return B2_ERR_CUDA;
  }
  std::vector<int> devs;
  if (!devices || n_devices <= 0) {
    for (int d = 0; d < ndev; ++d) devs.push_back(d);  // every device of the box
  } else {
    for (int i = 0; i < n_devices; ++i) {
      if (devices[i] < 0 || devices[i] >= ndev || std::count(devs.begin(), devs.end(), devices[i])) {
        b2::set_last_error("b2_group_create: device list entry %d (= %d) is outside [0, %d) or repeated", i,
                           devices[i], ndev);
        return B2_ERR_ARG;
      }
      devs.push_back(devices[i]);
    }
  }
  if (!cfg->model_path) {
    b2::set_last_error("b2_group_create: a group needs a model (b2_config.model_path)");
    return B2_ERR_ARG;
  }
  const int W = static_cast<int>(devs.size());
  b2_group* g = new b2_group();
  g->devices = devs;
  g->members.assign(W, nullptr);
  g->d_vec.assign(W, nullptr); g->d_ms.assign(W, nullptr);
  g->d_text.assign(W, nullptr); g->d_text_cap.assign(W, 0);
  g->d_off.assign(W, nullptr); g->d_off_cap.assign(W, 0);
  int rc = B2_OK;
  if (W > 1 && (rc = load_nccl(&g->nccl))) { delete g; return rc; }

  // members load their weights side by side (seconds each for BERT-base)
  std::vector<int> rcs(W, B2_OK);
  std::vector<std::string> errs(W);
  std::vector<std::thread> th;
  for (int i = 0; i < W; ++i) {
    th.emplace_back([&, i]() {
      b2_config c = *cfg;
      c.struct_size = sizeof(b2_config);
      c.device = devs[i];
      rcs[i] = b2_create(&c, &g->members[i]);
      if (rcs[i]) errs[i] = b2::last_error();
    });
  }
  for (auto& t : th) t.join();
  for (int i = 0; i < W && !rc; ++i) {
    if (rcs[i]) {
      b2::set_last_error("b2_group_create: device %d: %s", devs[i], errs[i].c_str());
      rc = rcs[i];
    }
  }
  if (!rc) {
    g->num_labels = b2::handle_num_labels(g->members[0]);
    if (W > 1) {
      g->comms.assign(W, nullptr);
      const ncclResult_t nr = g->nccl->CommInitAll(g->comms.data(), W, devs.data());
      if (nr != ncclSuccess) {
        b2::set_last_error("b2_group_create: ncclCommInitAll over %d devices failed: %s", W,
                           g->nccl->GetErrorString(nr));
        g->comms.clear();
        rc = B2_ERR_CUDA;
      }
      if (!rc && g->nccl->GetVersion) g->nccl->GetVersion(&g->nccl_version);
    }
  }
  if (!rc) {
    int prev = -1;
    cudaGetDevice(&prev);
    for (int i = 0; i < W && !rc; ++i) {
      cudaError_t e = cudaSetDevice(devs[i]);
      if (e == cudaSuccess) e = cudaMalloc(&g->d_vec[i], size_t(g->num_labels + N_EXTRA) * 8);
      if (e == cudaSuccess) e = cudaMalloc(&g->d_ms[i], 8);
      // direct loads/stores and copies between the members (the CSV chain sends a slice's text from
      // the device that indexed the file to the one that classifies it)
      for (int j = 0; j < W && e == cudaSuccess; ++j) {
        int can = 0;
        if (j != i && cudaDeviceCanAccessPeer(&can, devs[i], devs[j]) == cudaSuccess && can) {
          const cudaError_t pe = cudaDeviceEnablePeerAccess(devs[j], 0);
          if (pe != cudaSuccess) cudaGetLastError();  // already enabled is fine; copies fall back to staging
        }
      }
      if (e != cudaSuccess) {
        b2::set_last_error("b2_group_create: device %d: %s", devs[i], cudaGetErrorString(e));
        rc = B2_ERR_CUDA;
      }
    }
    if (prev >= 0) cudaSetDevice(prev);
  }
  if (rc) {
    const std::string keep = b2::last_error();
    b2_group_destroy(g);
    b2::set_last_error("%s", keep.c_str());
    return rc;
  }
  *out = g;
  return B2_OK;
}

void b2_group_destroy(b2_group* g) {
  if (!g) return;
  int prev = -1;
  cudaGetDevice(&prev);
  for (size_t i = 0; i < g->members.size(); ++i) {
    if (cudaSetDevice(g->devices[i]) != cudaSuccess) continue;
    if (i < g->comms.size() && g->comms[i]) g->nccl->CommDestroy(g->comms[i]);
    if (g->d_vec[i]) cudaFree(g->d_vec[i]);
    if (g->d_ms[i]) cudaFree(g->d_ms[i]);
    if (g->d_text[i]) cudaFree(g->d_text[i]);
    if (g->d_off[i]) cudaFree(g->d_off[i]);
  }
  if (prev >= 0) cudaSetDevice(prev);
  for (b2_handle* h : g->members) b2_destroy(h);
  delete g;
}

int32_t b2_group_size(b2_group* g) { return g ? static_cast<int32_t>(g->members.size()) : 0; }

b2_handle* b2_group_member(b2_group* g, int32_t i) {
  if (!g || i < 0 || i >= static_cast<int32_t>(g->members.size())) return nullptr;
  return g->members[i];
}

int b2_group_info(b2_group* g, int32_t* devices, int32_t max_devices, int32_t* nccl_version, int64_t* allreduces) {
  if (!g) { b2::set_last_error("b2_group_info: NULL group"); return B2_ERR_ARG; }
  for (int i = 0; devices && i < max_devices && i < static_cast<int>(g->devices.size()); ++i)
    devices[i] = g->devices[i];
  if (nccl_version) *nccl_version = g->nccl_version;
  if (allreduces) *allreduces = g->allreduces;
  return B2_OK;
}

int b2_group_classify_text(b2_group* g, const uint8_t* utf8, const int64_t* offsets, int64_t n, int32_t max_len,
                           int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                           int64_t* histogram, float* logits, b2_metrics* metrics, b2_metrics* per_device) {
  int rc;
  if ((rc = check_outputs(g, n, topk, top_idx, top_score))) return rc;
  if (!offsets || offsets[0] != 0 || (n > 0 && offsets[n] > 0 && !utf8)) {
    b2::set_last_error("b2_group_classify_text: offsets must start at 0 and utf8 must not be NULL");
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g->mu);
  CallerDevice restore;
  const int W = static_cast<int>(g->members.size()), C = g->num_labels;
  static const uint8_t empty = 0;
  const uint8_t* text = utf8 ? utf8 : &empty;
  std::vector<SliceResult> res(W);
  std::vector<int64_t> total;
  double max_ms = 0.0;
  auto work = [&](int i) {
    SliceResult& r = res[i];
    memset(&r.m, 0, sizeof(r.m));
    r.m.struct_size = sizeof(r.m);
    r.hist.assign(C, 0);
    if (cudaSetDevice(g->devices[i]) != cudaSuccess) {
      b2::set_last_error("cudaSetDevice(%d) failed", g->devices[i]);
      fail(&r, B2_ERR_CUDA);
    } else {
      const int64_t lo = i * n / W, hi = (i + 1) * n / W, cnt = hi - lo;  // contiguous slice
      std::vector<int64_t> off(cnt + 1);
      for (int64_t k = 0; k <= cnt; ++k) off[k] = offsets[lo + k] - offsets[lo];
      const int rc_i = b2_classify_text(g->members[i], text + offsets[lo], off.data(), cnt, max_len, topk,
                                        top_idx + lo * topk, top_score + lo * topk, labels ? labels + lo : nullptr,
                                        r.hist.data(), logits ? logits + lo * C : nullptr, &r.m);
      if (rc_i) fail(&r, rc_i);
    }
    reduce_member(g, i, &r, &total, &max_ms);
  };
  std::vector<std::thread> th;
  for (int i = 1; i < W; ++i) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  return finish_job(g, res, total, max_ms, histogram, metrics, per_device);
}

int b2_group_csv_classify(b2_group* g, b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows,
                          int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                          int64_t* histogram, float* logits, b2_metrics* metrics, b2_metrics* per_device) {
  int rc;
  if ((rc = check_outputs(g, n_rows, topk, top_idx, top_score))) return rc;
  if (!c || b2::csv_owner(c) != g->members[0]) {
    b2::set_last_error("b2_group_csv_classify: the CSV must have been opened on b2_group_member(group, 0)");
    return B2_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g->mu);
  CallerDevice restore;
  const int W = static_cast<int>(g->members.size()), C = g->num_labels;
  std::vector<SliceResult> res(W);
  std::vector<int64_t> total;
  double max_ms = 0.0;
  // Member 0 holds the file and its index. It cuts the text column of every slice out of the file
  // and sends it to the member that will classify it, device to device; the extracted text never
  // visits the host. Member 0 works through the slices in order 1, 2, ..., W-1, 0, and every member
  // starts on its slice the moment it has arrived.
  std::vector<std::thread> th;
  auto classify_slice = [&](int i, int64_t lo, int64_t cnt) {
    SliceResult& r = res[i];
    if (r.rc == B2_OK) {
      if (cudaSetDevice(g->devices[i]) != cudaSuccess) {
        b2::set_last_error("cudaSetDevice(%d) failed", g->devices[i]);
        fail(&r, B2_ERR_CUDA);
      } else {
        const int rc_i = b2::classify_resident_text(g->members[i], g->d_text[i], g->d_off[i], cnt, max_len, topk,
                                                    top_idx + lo * topk, top_score + lo * topk,
                                                    labels ? labels + lo : nullptr, r.hist.data(),
                                                    logits ? logits + lo * C : nullptr, &r.m);
        if (rc_i) fail(&r, rc_i);
      }
    }
    reduce_member(g, i, &r, &total, &max_ms);
  };
  int64_t lo0 = 0, cnt0 = 0;
  for (int step = 0; step < W; ++step) {
    const int i = (step + 1) % W;
    const int64_t lo = i * n_rows / W, hi = (i + 1) * n_rows / W, cnt = hi - lo;
    SliceResult& r = res[i];
    memset(&r.m, 0, sizeof(r.m));
    r.m.struct_size = sizeof(r.m);
    r.hist.assign(C, 0);
    int64_t bytes = 0;
    int rc_i = b2::csv_column_extract(c, column, start_row + lo, cnt, &bytes);
    if (!rc_i) {
      cudaSetDevice(g->devices[i]);
      rc_i = grow_dev(reinterpret_cast<void**>(&g->d_text[i]), &g->d_text_cap[i], size_t(bytes) + 64);
      if (!rc_i) rc_i = grow_dev(reinterpret_cast<void**>(&g->d_off[i]), &g->d_off_cap[i], size_t(cnt + 2) * 8);
      if (!rc_i) rc_i = b2::csv_column_copy_to(c, g->devices[i], g->d_text[i], g->d_off[i], cnt, bytes);
    }
    if (rc_i) fail(&r, rc_i);
    if (i == 0) { lo0 = lo; cnt0 = cnt; }
    else th.emplace_back(classify_slice, i, lo, cnt);
  }
  classify_slice(0, lo0, cnt0);
  for (auto& t : th) t.join();
  return finish_job(g, res, total, max_ms, histogram, metrics, per_device);
}

}  // extern "C"
