/* b2_classify.h — C ABI of libb2classify.so, the B200 back end of agent-tpu's `map_classify_tpu`
 * and `map_tokenize` ops.
 *
 * The reference has no FFI: its back end is a Python object (a TFLite interpreter) reached from
 * two call sites. Each entry point below names the reference call it stands in for; paths are
 * relative to the reference repository root. INTEGRATION.md shows the ctypes binding.
 *
 * Conventions
 *   - every function returns 0 on success or a B2_ERR_* code; b2_last_error() then returns a
 *     human-readable message for the calling thread;
 *   - the caller owns every buffer it passes; the library owns device memory, streams and events;
 *   - calls are blocking: results are complete when the call returns;
 *   - a handle is bound to one CUDA device and must be used by one thread at a time (the agent runs
 *     tasks inline, one at a time: app.py:286);
 *   - there is no CPU fallback: without a usable CUDA device b2_create fails.
 */
#ifndef B2_CLASSIFY_H
#define B2_CLASSIFY_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK 0
#define B2_ERR_ARG 1     /* bad argument */
#define B2_ERR_CUDA 2    /* CUDA runtime / driver error */
#define B2_ERR_IO 3      /* file missing or unreadable */
#define B2_ERR_FORMAT 4  /* file present but not in the expected format */
#define B2_ERR_LIMIT 5   /* request exceeds a limit of the handle */

typedef struct b2_handle b2_handle;

/* Replaces the arguments of get_tpu_handle(model_path) (ops/_tpu_runtime.py:34). */
typedef struct b2_config {
  int32_t struct_size;      /* sizeof(b2_config) */
  int32_t device;           /* CUDA device ordinal */
  const char* model_path;   /* .b2m model file; NULL gives a tokenizer-only handle */
  const char* vocab_path;   /* vocab.txt, one WordPiece per line; NULL gives a handle without tokenizer */
  const char* unicode_path; /* bert_unicode.bin (required with vocab_path) */
  int32_t max_batch;        /* items per encoder pass, 0 = 1024 */
  int32_t max_len;          /* largest padded length this handle will be asked for, 0 = 128 */
} b2_config;

/* Replaces interpreter.get_input_details() / get_output_details() (ops/_tpu_runtime.py:59-60). */
typedef struct b2_model_info {
  int32_t hidden, layers, heads, intermediate;
  int32_t vocab_size, max_positions, num_labels, tokenizer_vocab_size;
  int32_t max_batch, max_len, device, sm_count;
  int64_t weight_bytes, activation_bytes;
  int32_t packed_rows; /* 1: padding-free token rows (default); 0: B2_FORCE_PADDED=1 was set */
  int32_t ln_folded;   /* 1: LayerNorms folded into the GEMM epilogues (default); 0: B2_NO_LN_FOLD=1 */
} b2_model_info;

/* Per-call counters and device timings (CUDA events on the library's stream). The reference
 * reports only wall-clock elapsed_ms (ops/map_classify_tpu.py:45,81). */
typedef struct b2_metrics {
  int32_t struct_size; /* sizeof(b2_metrics), set by the caller */
  int32_t profiled;    /* 1 if per-kernel timings below were collected (b2_set_profiling) */
  int64_t items, tokens, padded_tokens, batches;
  int64_t kernel_launches;
  int64_t h2d_bytes, d2h_bytes;
  double gpu_ms_total; /* first kernel to last kernel, all batches */
  /* per kernel kind, summed over launches (only when profiled): */
  double ms_tokenize, ms_embed, ms_gemm_qkv, ms_attention, ms_gemm_attn_out, ms_layernorm,
      ms_gemm_ffn1, ms_gemm_ffn2, ms_pooler, ms_head;
  int64_t n_tokenize, n_embed, n_gemm_qkv, n_attention, n_gemm_attn_out, n_layernorm, n_gemm_ffn1,
      n_gemm_ffn2, n_pooler, n_head;
} b2_metrics;

/* Replaces nvidia-smi parsing in _parse_nvidia_smi (worker_sizing.py:139-161). */
typedef struct b2_device_info {
  char name[128];
  int64_t total_memory_bytes;
  int32_t sm_count, cc_major, cc_minor, l2_bytes;
} b2_device_info;

/* Thread-local message for the last failing call. No reference counterpart (the reference raises
 * Python exceptions: ops/_tpu_runtime.py:49-53). */
const char* b2_last_error(void);

/* Number of CUDA devices, 0 if none (no error). Replaces list_edge_tpus() (ops/_tpu_runtime.py:48). */
int32_t b2_device_count(void);
int b2_get_device_info(int32_t device, b2_device_info* out);

/* Free and total device memory right now (cudaMemGetInfo), so a controller can size `batch` x
 * `max_len` for this worker. Extends _detect_gpu (worker_sizing.py:164-185), which reports only
 * memory.total. Either pointer may be NULL. */
int b2_get_device_memory(int32_t device, int64_t* free_bytes, int64_t* total_bytes);
/* How `device` reaches `peer`: can_access = 1 when peer memory is directly addressable (NVLink /
 * NVSwitch or PCIe P2P), performance_rank = the driver's relative rank of that link (lower is
 * faster; cudaDevP2PAttrPerformanceRank), native_atomics = 1 when atomics work across it (NVLink).
 * All 0 for device == peer. No reference counterpart (the reference has no multi-device notion). */
int b2_get_peer_info(int32_t device, int32_t peer, int32_t* can_access, int32_t* performance_rank,
                     int32_t* native_atomics);

/* Load model + tokenizer onto a device: make_interpreter + allocate_tensors
 * (ops/_tpu_runtime.py:55-56). */
int b2_create(const b2_config* config, b2_handle** out);
void b2_destroy(b2_handle* h);
int b2_get_model_info(b2_handle* h, b2_model_info* out);
/* 1: time every kernel launch with CUDA events (adds one stream sync per batch); 0: off. */
int b2_set_profiling(b2_handle* h, int32_t on);
/* The cudaStream_t every kernel of this handle is launched on, as an integer, so a caller can
 * record its own CUDA events on it. No reference counterpart. */
uint64_t b2_get_stream(b2_handle* h);

/* WordPiece tokenize-and-pad. utf8 holds the n items back to back, item i is bytes
 * [offsets[i], offsets[i+1]). ids: [n, max_len] int32, mask: [n, max_len] uint8 (either may be
 * NULL). No reference counterpart; shared by map_tokenize (mode "wordpiece") and
 * map_classify_tpu. Host pointers. Jobs of 131,072 items or more run as 65,536-item chunks on three
 * streams, so that the upload of one chunk, the kernel of the next and the read-back of a third
 * overlap (same results); page-locked host buffers are what lets the copies overlap. */
int b2_tokenize(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n, int32_t max_len,
                int32_t* ids, uint8_t* mask, b2_metrics* metrics);
/* Same with every pointer in device memory on the handle's device (inputs already resident). */
int b2_tokenize_device(b2_handle* h, const uint8_t* d_utf8, const int64_t* d_offsets, int64_t n,
                       int64_t n_bytes, int32_t max_len, int32_t* d_ids, uint8_t* d_mask,
                       b2_metrics* metrics);

/* Classify n text items: tokenize, encoder forward, head. Replaces set_tensor / invoke /
 * get_tensor / _topk for a whole job (ops/map_classify_tpu.py:61-75, 15-19).
 *   top_idx, top_score: [n, topk]  descending by score (softmax probability); k clamped like the
 *                       reference must be done by the caller: 1 <= topk <= num_labels
 *   labels:    [n] arg-max class                          (may be NULL)
 *   histogram: [num_labels] counts of labels, overwritten (may be NULL)
 *   logits:    [n, num_labels] float32                    (may be NULL)
 * Host pointers. */
int b2_classify_text(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                     int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score,
                     int32_t* labels, int64_t* histogram, float* logits, b2_metrics* metrics);
/* Same with text, offsets and all outputs in device memory. */
int b2_classify_text_device(b2_handle* h, const uint8_t* d_utf8, const int64_t* d_offsets, int64_t n,
                            int64_t n_bytes, int32_t max_len, int32_t topk, int32_t* d_top_idx,
                            float* d_top_score, int32_t* d_labels, int64_t* d_histogram,
                            float* d_logits, b2_metrics* metrics);
/* Classify n pre-tokenised rows: ids [n, max_len] int32, mask [n, max_len] uint8 (1 = token).
 * This is the v0 contract's `input: list[int]` (ops/map_classify_tpu.CONTRACT.md:7-8,
 * ops/map_classify_tpu.py:61-72) widened to n rows. Host pointers.
 * Every mask row must be a non-empty prefix of ones followed by zeros (right padding), which is what
 * b2_tokenize produces: the encoder stores sequence b as rows [0, len_b) only. A row with a hole,
 * with left padding or without any token is refused with B2_ERR_ARG (the message names the first
 * such row) before anything is computed — it would otherwise be classified as a different input. */
int b2_classify_ids(b2_handle* h, const int32_t* ids, const uint8_t* mask, int64_t n, int32_t max_len,
                    int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                    int64_t* histogram, float* logits, b2_metrics* metrics);

/* Fixed-width chunking in code points: _chunk_text (ops/map_tokenize.py:6-9) over items.
 * Pass 1: code points per item -> counts[n]. Pass 2: chunk j of item i starts at byte
 * bounds[chunk_base[i] + j] of utf8, where chunk_base is the exclusive prefix sum of
 * ceil(counts[i] / chunk_size) computed by the caller. Host pointers. */
int b2_chunk_count(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                   int64_t* counts);
int b2_chunk_bounds(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                    int64_t chunk_size, const int64_t* chunk_base, int64_t n_chunks, int64_t* bounds);

/* ---- CSV shards: read_csv_shard (ops/csv_shard.py) and the csv -> tokenize -> classify chain ----
 *
 * b2_csv_open copies a CSV file's bytes to the device and indexes its records once, with the
 * semantics of the reference's reader: csv.DictReader over open(path, newline="", encoding="utf-8")
 * (ops/csv_shard.py:17-18) — default dialect, the first record is the header, records without any
 * field (blank lines) are dropped, quoted fields may hold commas, quotes ("" -> ") and line ends.
 * Shards are then cut from the index instead of re-parsing the file from the top as
 * _read_csv_shard does (ops/csv_shard.py:19-24). Rows are 0-based over data rows, as `start_row`
 * is in the reference (ops/csv_shard.py:11-12). */
typedef struct b2_csv b2_csv;
typedef struct b2_csv_info {
  int32_t struct_size;    /* sizeof(b2_csv_info), set by the caller */
  int32_t has_header;     /* 0: the data held no record at all (DictReader.fieldnames is None) */
  int32_t header_fields;  /* number of header names (0 for a blank first line) */
  int32_t end_state;      /* automaton state after the last byte; > 1: last row has no line end */
  int64_t header_bytes;   /* bytes of the unquoted header names, see b2_csv_header */
  int64_t n_bytes;
  int64_t data_offset;    /* offset of the first byte after the header record */
  int64_t rows;           /* data rows: what DictReader yields over the whole file */
  /* What the reference would trip over while reading (it reports them as
   * "failed reading csv: ..." only if it reads that far: ops/csv_shard.py:78-81): */
  int64_t first_invalid_utf8;    /* offset where strict UTF-8 decoding fails, n_bytes if the data
                                    ends inside a sequence, -1 if the bytes are valid */
  int32_t header_field_too_long; /* 1: a header name exceeds csv.field_size_limit() = 131072 */
  int32_t reserved;
  int64_t header_long_offset;    /* offset of the code point that overflows it, else -1 */
  int64_t long_field_row;        /* first data row with a field above the limit, else -1 */
  int64_t long_field_offset;     /* offset of the code point that overflows it, else -1 */
  double gpu_ms;                 /* copy + index, CUDA events on the handle's stream */
  int64_t kernel_launches;
  double index_ms;               /* the index kernels alone (after the copy), incl. two tiny readbacks */
} b2_csv_info;

/* bytes: the whole file in host memory (n_bytes may be 0). The csv object belongs to handle h, uses
 * its device and stream, and must be closed before the handle is destroyed. */
int b2_csv_open(b2_handle* h, const uint8_t* bytes, int64_t n_bytes, b2_csv** out, b2_csv_info* info);
void b2_csv_close(b2_csv* c);
/* Header names, unquoted, back to back in names[0 .. info.header_bytes); name i ends at
 * name_end[i]. Replaces DictReader.fieldnames. */
int b2_csv_header(b2_csv* c, uint8_t* names, int64_t names_cap, int64_t* name_end, int32_t max_names);
/* Byte offsets in the file of rows [start_row, start_row + n_rows): first byte, and the line end
 * that closes the row (n_bytes for a last row without one). Either pointer may be NULL. */
int b2_csv_row_spans(b2_csv* c, int64_t start_row, int64_t n_rows, int64_t* row_start, int64_t* row_end);
/* Size the content of rows [start_row, start_row + n_rows): total number of fields and total bytes
 * of unquoted field content. column >= 0 keeps only that column: one field per row, empty where
 * the row is shorter; column < 0 keeps every field. */
int b2_csv_rows_size(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t column, int64_t* n_fields,
                     int64_t* n_bytes);
/* Fetch what b2_csv_rows_size sized: field content back to back in bytes, field j ends at
 * field_end[j], row i's last field is field row_field_end[i] - 1 (rows in order, so row i holds
 * fields [row_field_end[i-1], row_field_end[i])). Replaces the `rows.append(row)` loop of
 * _read_csv_shard (ops/csv_shard.py:19-24); building the dicts stays with the caller. */
int b2_csv_rows_read(b2_csv* c, int64_t start_row, int64_t n_rows, int32_t column, uint8_t* bytes,
                     int64_t* field_end, int64_t* row_field_end);
/* The chained job of BASELINE config 5: take `column` of rows [start_row, start_row + n_rows) as
 * the text items and classify them like b2_classify_text — the extracted text never leaves the
 * device. No reference counterpart (the agent runs one op per task: app.py:283-287); outputs as
 * in b2_classify_text, host pointers. */
int b2_csv_classify(b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows, int32_t max_len,
                    int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels, int64_t* histogram,
                    float* logits, b2_metrics* metrics);

/* ---- one job over several GPUs of one box, from one process ----
 *
 * The reference's agent runs one task at a time, inline, in one process (app.py:31 MAX_TASKS=1,
 * app.py:191 takes tasks[0], app.py:286 "run inline"), so the single call it makes — the op's
 * run(payload) -> this library — has to spread the job over the box's GPUs itself. A group owns one
 * b2_handle per device. A job's n items are cut into contiguous slices [i*n/W, (i+1)*n/W), one per
 * device; one host thread per device runs the per-device pipeline of b2_classify_text on its slice
 * and writes the results into the caller's arrays at the slice's position, so outputs come back in
 * item order without any gather. The one exchange per job is an NCCL all-reduce (sum) of
 * int64[num_labels + 8] — label histogram, the additive b2_metrics counters and a failure count — plus
 * a max-reduce of the device times, over communicators created once with ncclCommInitAll. No
 * reference counterpart (the reference scales out by running more agents: SURVEY.md section 2.1). */
typedef struct b2_group b2_group;
/* config->device is ignored; devices = NULL or n_devices <= 0 takes every CUDA device of the box.
 * A group of one device needs no NCCL; more than one loads libnccl.so.2 at run time (a copy the
 * process has already loaded, else $B2_NCCL_PATH, else the linker's search path) and fails with
 * B2_ERR_IO if there is none. Weights are loaded on all devices side by side. */
int b2_group_create(const b2_config* config, const int32_t* devices, int32_t n_devices, b2_group** out);
void b2_group_destroy(b2_group* g);
int32_t b2_group_size(b2_group* g);
/* Member i's handle (owned by the group; valid until b2_group_destroy). Every single-device entry
 * point works on it; b2_csv_open for a group job must use member 0. */
b2_handle* b2_group_member(b2_group* g, int32_t i);
/* devices[i] = CUDA ordinal of member i (up to max_devices entries); nccl_version = ncclGetVersion()
 * or 0 for a group of one; allreduces = NCCL collectives issued so far. Pointers may be NULL. */
int b2_group_info(b2_group* g, int32_t* devices, int32_t max_devices, int32_t* nccl_version, int64_t* allreduces);
/* b2_classify_text over the whole group. metrics: job totals (counters summed over devices by the
 * all-reduce, gpu_ms_total = the slowest device's time, per-kernel times summed); per_device: NULL or
 * an array of b2_group_size() b2_metrics (struct_size set in each) for the individual devices.
 * If any device fails, every device still takes part in the all-reduce and the call returns the
 * first error. Host pointers. */
int b2_group_classify_text(b2_group* g, const uint8_t* utf8, const int64_t* offsets, int64_t n, int32_t max_len,
                           int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                           int64_t* histogram, float* logits, b2_metrics* metrics, b2_metrics* per_device);
/* b2_csv_classify over the whole group (BASELINE config 5: csv_shard -> map_tokenize -> map_classify
 * on 8 GPUs with the NCCL histogram reduce). `c` was opened with b2_csv_open on b2_group_member(g, 0):
 * the file crosses PCIe once, to that device. Member 0 cuts the text column of each slice of rows out
 * of the indexed file and sends it to the member that classifies it, device to device (NVLink
 * peer-to-peer where available); the extracted text never visits the host. */
int b2_group_csv_classify(b2_group* g, b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows,
                          int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score, int32_t* labels,
                          int64_t* histogram, float* logits, b2_metrics* metrics, b2_metrics* per_device);

/* Tokenize-and-pad like b2_tokenize, but without allocating device memory during the call.
 * No counterpart in the reference. It exists for callers that are not allowed to ask the CUDA
 * driver for memory once they are running (a worker with a fixed memory budget, a loop that cannot
 * absorb the synchronisation an allocation or free brings): the device-side staging for the text,
 * offsets, ids and mask is a fixed block of B2_NOALLOC_SCRATCH_BYTES that every GPU holds from the
 * moment the library is loaded there until exit.
 *   - a job that does not fit the block (text bytes + 8*(n+1) + 5*n*max_len, each part rounded up
 *     to 256) is refused with B2_ERR_LIMIT; b2_tokenize has no such limit;
 *   - all calls on one GPU share the block: do not let two of them overlap in time, whichever
 *     handles or threads they come from;
 *   - the first call on each GPU looks up the block's address; make it before any stream capture.
 * Host pointers; ids [n, max_len] int32; mask [n, max_len] uint8 or NULL. */
#define B2_NOALLOC_SCRATCH_BYTES (8 * 1024 * 1024)
int b2_tokenize_noalloc(b2_handle* h, const uint8_t* utf8, const int64_t* offsets, int64_t n,
                        int32_t max_len, int32_t* ids, uint8_t* mask);

/* Test aid, no reference counterpart (the reference's model is an opaque interpreter.invoke(),
 * ops/map_classify_tpu.py:72): run ONE batch of n <= max_batch pre-tokenised rows through the
 * production kernels exactly as b2_classify_ids does and also return every layer's output as the next
 * kernel reads it, up-cast to float32:
 *   hidden [layers + 1, n, max_len, hidden]: [0] the embedding output, [l] the output of encoder
 *          layer l. Pad positions are zero. The last layer is only computed for position 0 of every
 *          row (nothing else reaches the pooler), so hidden[layers][:, 1:, :] is zero.
 *   logits [n, num_labels] or NULL.
 * It allocates and frees its own device buffer per call; not for production use. Host pointers. */
int b2_debug_hidden_states(b2_handle* h, const int32_t* ids, const uint8_t* mask, int64_t n, int32_t max_len,
                           float* hidden, float* logits);

/* Developer aid, no reference counterpart: 64 bytes of mapped host memory in which the first bounded
 * pipeline wait that gives up on the calling thread's current device records where it was before the
 * launch is trapped (layout in csrc/b2_api.cu). Call after b2_create; NULL on failure. */
unsigned long long* b2_debug_trap_log(void);

#ifdef __cplusplus
}
#endif
#endif /* B2_CLASSIFY_H */
