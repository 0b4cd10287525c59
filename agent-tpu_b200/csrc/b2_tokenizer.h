// b2_tokenizer.h — device-resident tokenizer state and launchers (b2_tokenizer.cu).
#pragma once
#include "b2_kernels.h"

namespace b2 {

struct VocabSlot {
  uint64_t key;      // 0 = empty
  uint32_t id;       // token id (line number in vocab.txt)
  uint32_t off_len;  // (offset into vocab_cps << 8) | length in code points
};

// Passed to the kernel by value: table pointers live in device memory.
struct TokenizerDev {
  const uint16_t* page_index;  // code point >> 7 -> unique page
  const uint32_t* pages;       // per-code-point normalisation entries (gen_unicode_tables.py)
  const uint32_t* pool;        // multi-code-point expansions
  const VocabSlot* table;      // open-addressing hash of the vocabulary
  const uint32_t* vocab_cps;   // code points of every vocabulary piece (without "##")
  const uint64_t* pows;        // HASH_P^k, k = 0..100
  int table_bits;
  int unk_id, cls_id, sep_id, pad_id;
  int n_special;               // literal special tokens recognised in raw text
  int special_len[5];
  int special_id[5];
  char special_text[5][8];
};

int tokenizer_load(const char* vocab_path, const char* unicode_path, TokenizerDev* T, int* vocab_size);
void tokenizer_free(TokenizerDev* T);
// text: concatenated UTF-8 of all lines; offsets[n_lines + 1]: byte range of line i.
// ids [n_lines, max_len] i32, mask [n_lines, max_len] u8 (may be null), lens [n_lines] i32 (may be null)
int tokenizer_run(cudaStream_t stream, const TokenizerDev& T, const uint8_t* text,
                  const int64_t* offsets, int n_lines, int max_len, int32_t* ids, uint8_t* mask,
                  int32_t* lens, int num_sms);
// code points per item / byte offsets of every `chunk_size`-th code point (ops/map_tokenize.py:6-9)
int chunk_count(cudaStream_t stream, const uint8_t* text, const int64_t* offsets, int n,
                int64_t* counts);
int chunk_bounds(cudaStream_t stream, const uint8_t* text, const int64_t* offsets, int n,
                 int64_t chunk_size, const int64_t* chunk_base, int64_t* bounds);

}  // namespace b2
