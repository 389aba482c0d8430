// Host interface of the GEMM-based T3 prompt prefill (csrc/t3_prefill.cu), called by cbx_t3_prefill.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace cbx {

struct T3PrefillArgs {
  const float* embeds;       // [B][P][1024] fp32 (t3.py:228-252)
  int B, P, n_layers, max_pos, kv_fp32;
  const void* wplain;        // bf16, per layer [q|k|v 3072x1024][o 1024x1024][gate/up 8192x1024, 128 gate + 128 up rows per
                             // 256-row tile][down 1024x4096], then speech_head [8194][1024]
  const float *ln1, *ln2, *final_norm, *rope_cos, *rope_sin;
  void *kcache, *vcache;     // the decode kernel's slab [layer][2][16][max_pos][64]
  float* logits;             // [2][logits_ld] pending logits of the last prompt position
  int logits_ld;             // floats between the two rows (0 = 8194)
  int* state;                // {pos, n_gen, stop, err}
  void* workspace;           // t3_prefill_workspace_bytes(max_pos)
  // several utterances of the SAME prompt length in one pass (batched engine): B = 2 nseq rows, utterance u = rows 2u, 2u + 1;
  // its KV slab starts kv_seq_stride ELEMENTS after utterance u - 1's, its logits rows 2 logits_ld floats after.  nseq <= 1: one
  // utterance (B = 1 or 2), the fields below unused.  ws_rows: row capacity of `workspace` (0 = 2 max_pos).
  int nseq;
  long long kv_seq_stride;
  long long ws_rows;
  // alignment analyzer (al_layer < 0: off): head-mean attention of CFG row 0 over text columns [al_i, al_i + al_n) for the
  // last al_rows0 prompt positions at layer al_layer -> al_out [al_rows0][al_cap]   (alignment_stream_analyzer.py:96-101)
  int al_layer, al_i, al_n, al_rows0, al_cap;
  float* al_out;
};

size_t t3_prefill_workspace_bytes(int max_pos);
size_t t3_prefill_workspace_bytes_rows(size_t rows);       // for `rows` activation rows (batched prefill)
int t3_prefill_gemm(const T3PrefillArgs& a, cudaStream_t st);

}  // namespace cbx
