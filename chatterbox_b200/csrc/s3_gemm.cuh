// Host-visible interface of the tcgen05 GEMM used by the S3Gen estimator (csrc/s3_gemm.cu).
#pragma once
#include "tc.cuh"

namespace cbx {

struct GemmEpi {
  const float* bias;                       // [N] or null
  int bias_mod;                            // > 0: bias index is n % bias_mod (polyphase transposed convolution)
  const float* ln_g; const float* ln_b;    // LayerNorm over the N outputs of a row, then Mish (needs n_tile == N)
  const float* temb;                       // [N] added after Mish (the same time embedding for every batch row), or null
  int gelu;                                // exact (erf) GELU
  int elu;                                 // ELU(alpha = 1)
  const float* resid; int ld_resid;        // fp32 [rows][ld] added last
  const float* resid2; int ld_resid2;      // second fp32 addend (running sum over parallel branches)
  long long out_shift, out_limit;          // out_f32 only: flat element offset added to row*ld + n; stores outside
                                           // [0, out_limit) are dropped (out_limit 0 = no shift / no clipping)
  const float* snake_alpha;                // bf16 output = v + sin^2(alpha[n] v) / (alpha[n] + 1e-9)   (HiFT Snake)
  float lrelu;                             // bf16 output = leaky_relu(v, lrelu) when != 0
  float* out_f32; int ld_f32;
  __nv_bfloat16* out_bf16; int ld_bf16;
  const float* ln2_g; const float* ln2_b;  // LayerNorm of the final row value -> out_ln (needs out_f32, n_tile == N)
  __nv_bfloat16* out_ln; int ld_ln;
  // T3 prefill: the value leaves as TWO bf16 planes (hi = bf16(v), lo = bf16(v - hi)) so the next GEMM keeps ~16
  // mantissa bits of its activations; plane 1 starts hilo_plane elements after plane 0
  __nv_bfloat16* out_hilo; int ld_hilo; long long hilo_plane;
  int swiglu;                              // n_tile == 256 holds 128 gate then the 128 matching up columns:
                                           // v = silu(gate) * up, 128 output columns per tile (Llama MLP)
};

// C[b][t][n] = sum_tap sum_k A[b][t + tap*dil - pad][k] * W[n][tap*K + k]   (rows outside [0, L) read as zero),
// for t in [0, M_out)
struct GemmDesc {
  CUtensorMap tmA;     // bf16 [B][L][K] view (K innermost), box {64, 128, 1}
  CUtensorMap tmW;     // bf16 [N][taps*K], box {64, n_tile, 1}
  int B, L, N, K, taps, n_tile, dil, pad, M_out;
  int ksplit;          // CTAs per cluster sharing one output tile, each accumulating 1/ksplit of the K panels
  int nsplit;          // > 1: the N / n_tile column tiles of a row tile form one cluster; a row LayerNorm exchanges its
                       // per-row moments through distributed shared memory (no K split in this mode)
  int hilo;            // A is two planes (tensor-map dim 2): tap t reads plane t against the SAME W panel
  GemmEpi epi;
};

#ifdef __CUDACC__
// exact (erf) GELU as relu(x) - |x| h(|x|), h(a) = erfc(a / sqrt 2) / 2 = 2^P(a): P is a degree-6 least-squares fit of log2 h on
// [0, 6] (weighted by a h, the quantity that reaches the result); |x| is clamped at 6, where h = 1e-9.  Max |error| 2.5e-7 in
// fp32 over [-9, 9] — the Abramowitz-Stegun 7.1.26 form it replaces had 2.1e-7 — three orders below the bf16 rounding of the
// value this feeds.  One MUFU and 9 FMA-pipe instructions instead of two and 16: the up-projection epilogue of the padded-batch
// solve is issue-bound on this function (profiles/r02_z_gemm_batch_ncu.md).  Fit: numpy lstsq on 4000 Chebyshev nodes.
__device__ __forceinline__ float gelu_f(float x) {
  float a;                                   // min(|x|, 6) that keeps a NaN (fminf would turn it into 6 and hide it)
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(a) : "f"(fabsf(x)), "f"(6.0f));
  float p = fmaf(3.33725729433354e-05f, a, -0.0007683017174713314f);
  p = fmaf(p, a, 0.008067665621638298f);
  p = fmaf(p, a, -0.053377795964479446f);
  p = fmaf(p, a, -0.4588061571121216f);
  p = fmaf(p, a, -1.1511874198913574f);
  p = fmaf(p, a, -0.9999948143959045f);
  float h;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(h) : "f"(p));
  return fmaf(-a, h, fmaxf(x, 0.0f));          // a, not |x|: beyond the clamp the term is < 6e-9 either way, and +-inf stay +inf / 0
}
#endif

// builds the tensor maps; A has row stride lda elements (>= K), W is dense [N][taps*K]
int gemm_make(GemmDesc* d, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int B, int L, int N, int K, int taps,
              const GemmEpi& epi);    // causal taps: dil 1, pad taps-1, M_out = L
int gemm_make_conv(GemmDesc* d, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int B, int L, int N, int K, int taps,
                   int dil, int pad, int M_out, const GemmEpi& epi);
// C[m][n] = sum_k (A_hi[m][k] + A_lo[m][k]) * W[n][k]: A = two bf16 planes [2][M][lda], plane stride in elements
int gemm_make_hilo(GemmDesc* d, const __nv_bfloat16* A, int lda, long long plane, const __nv_bfloat16* W, int M, int N, int K,
                   const GemmEpi& epi);
// same with an explicit column tile (0 = widest divisor of N), explicit split-K cluster size (0 = automatic) and a
// weight matrix of N_w <= N rows (output columns >= N_w are zero): the skinny (M <= 128) decode GEMMs of the batched
// T3 engine want many narrow tiles so that every SM streams weights
int gemm_make_hilo_ex(GemmDesc* d, const __nv_bfloat16* A, int lda, long long plane, const __nv_bfloat16* W, int M, int N, int N_w,
                      int K, const GemmEpi& epi, int n_tile, int ksplit);
int gemm_launch(const GemmDesc& d, cudaStream_t st);
// true unless CBX_PDL=0: kernels are launched with programmatic stream serialization (prologue overlap)
bool pdl_enabled();

// ---- fused feed-forward of a BasicTransformerBlock (csrc/s3_ffn.cu; matcha/transformer.py:300-316):
//   y = resid + W2 gelu(W1 x + b1) + b2, then either LayerNorm(y) -> out_ln (the next block's norm1) or y -> out_bf16,
// x = bf16 [B][L][256] (already normalised), hidden 1024.  One cluster of 8 CTAs per 128-row tile: CTA c owns hidden
// columns [128 c, 128 c + 128) — FF1 slice, GELU, FF2 partial over that K slice — and the 8 partial tiles meet in
// distributed shared memory, each CTA finishing 16 rows.
struct FfnDesc {
  CUtensorMap tmX, tmW1, tmW2;
  int B, L;
  const float *b1, *b2;
  const float* resid; int ld_resid;
  float* out_f32; int ld_f32;
  const float *ln_g, *ln_b; __nv_bfloat16* out_ln; int ld_ln;      // ln_g == null: no LayerNorm output
  __nv_bfloat16* out_bf16; int ld_bf16;                            // or null
  // optional fused attention out-projection in front (ffn_make_out): x1 = resid0 + att Wo^T + bo -> out0_f32, and
  // LayerNorm(x1; g0, b0) -> the x tensor of ffn_make (the kernel then reads it back tile by tile)
  int fuse_out;
  CUtensorMap tmA, tmWo;
  const float *bo, *g0, *b0, *resid0; int ld_res0;
  float* out0_f32; int ld_out0;
  __nv_bfloat16* xbuf; int ld_x;
  float* red;          // optional scratch, (row tiles) x 8 x 128 x 256 fp32: cluster reduce through L2 instead of DSMEM
};
int ffn_make(FfnDesc* d, const __nv_bfloat16* X, int ldx, const __nv_bfloat16* W1, const __nv_bfloat16* W2, int B, int L);
// adds the out-projection: att = bf16 [B][L][512] (row stride lda), Wo = bf16 [256][512]; X of ffn_make must be d->xbuf
int ffn_make_out(FfnDesc* d, const __nv_bfloat16* att, int lda, const __nv_bfloat16* Wo);
int ffn_launch(const FfnDesc& d, cudaStream_t st);

}  // namespace cbx
