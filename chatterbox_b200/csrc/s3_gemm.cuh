// Host-visible interface of the tcgen05 GEMM used by the S3Gen estimator (csrc/s3_gemm.cu).
#pragma once
#include "tc.cuh"

namespace cbx {

struct GemmEpi {
  const float* bias;                       // [N] or null
  const float* ln_g; const float* ln_b;    // LayerNorm over the N outputs of a row, then Mish (needs n_tile == N)
  const float* temb;                       // [B][N] added after Mish, or null
  int gelu;                                // exact (erf) GELU
  const float* resid; int ld_resid;        // fp32 [rows][ld] added last
  float* out_f32; int ld_f32;
  __nv_bfloat16* out_bf16; int ld_bf16;
  const float* ln2_g; const float* ln2_b;  // LayerNorm of the final row value -> out_ln (needs out_f32, n_tile == N)
  __nv_bfloat16* out_ln; int ld_ln;
};

// C[b][t][n] = sum_tap sum_k A[b][t + tap - (taps-1)][k] * W[n][tap*K + k]   (rows t < 0 read as zero)
struct GemmDesc {
  CUtensorMap tmA;     // bf16 [B][L][K] view (K innermost), box {64, 128, 1}
  CUtensorMap tmW;     // bf16 [N][taps*K], box {64, n_tile, 1}
  int B, L, N, K, taps, n_tile;
  GemmEpi epi;
};

// builds the tensor maps; A has row stride lda elements (>= K), W is dense [N][taps*K]
int gemm_make(GemmDesc* d, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int B, int L, int N, int K, int taps,
              const GemmEpi& epi);
int gemm_launch(const GemmDesc& d, cudaStream_t st);

}  // namespace cbx
