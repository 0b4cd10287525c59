// b2_internal.h — what b2_group.cu needs from b2_api.cu beyond the exported C ABI. Not exported
// (C++ linkage inside namespace b2); the public surface stays include/b2_classify.h.
#pragma once
#include "b2_kernels.h"

struct b2_handle;
struct b2_csv;

namespace b2 {

int handle_device(const b2_handle* h);
int handle_num_labels(const b2_handle* h);
bool handle_has_model(const b2_handle* h);

// b2_classify_text for text that is already in the memory of h's device (d_text, d_off[n + 1] with
// d_off[0] == 0); outputs go to the caller's HOST arrays. Takes the handle's lock.
int classify_resident_text(b2_handle* h, const uint8_t* d_text, const int64_t* d_off, int64_t n,
                           int32_t max_len, int32_t topk, int32_t* top_idx, float* top_score,
                           int32_t* labels, int64_t* histogram, float* logits, b2_metrics* metrics);

// The handle an open CSV belongs to.
b2_handle* csv_owner(b2_csv* c);
// Extract `column` of rows [start_row, start_row + n_rows) on the CSV's own device (the tokenizer's
// utf8 / offsets pair) and report its size; the extraction stays valid until the next call on `c`.
int csv_column_extract(b2_csv* c, int32_t column, int64_t start_row, int64_t n_rows, int64_t* n_bytes);
// Copy the last extraction to device `dst_device`: n_bytes of text to d_text and n_rows + 1 offsets
// to d_off (peer-to-peer over NVLink when the devices can reach each other). Blocks until done.
int csv_column_copy_to(b2_csv* c, int dst_device, uint8_t* d_text, int64_t* d_off, int64_t n_rows,
                       int64_t n_bytes);

}  // namespace b2
