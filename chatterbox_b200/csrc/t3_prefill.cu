// T3 prompt prefill as batched GEMMs on the tcgen05 path (reference: the first `patched_model(inputs_embeds=...)`
// call of T3.inference, src/chatterbox/models/t3/t3.py:307-316 -> HF LlamaModel.forward over the whole prompt).
//
// The decode megakernel (t3_decode.cu) is a weight-streaming GEMV: run over a P-position prompt it streams the
// 1 GB of weights P times.  Here the prompt's 2 P rows (CFG pair) go through each matrix once:
//   x (fp32 residual stream, [B*P][1024])
//   -> k_rms_hilo: RMSNorm gain * x * rstd, written as TWO bf16 planes (hi, lo = bf16(v - hi))
//   -> q|k|v GEMM (A = hi plane + lo plane against the same bf16 weights: ~16 mantissa bits, fp32 accumulate in TMEM)
//   -> k_rope_kv: llama3 RoPE on q, k; K / V appended to the decode kernel's KV slab (same rounding as decode)
//   -> k_attn_prefill: causal softmax(q k^T / 8) v in fp32 on CUDA cores (42 MFLOP per layer at P = 102)
//   -> o_proj GEMM + residual, RMSNorm, gate/up GEMM with the SwiGLU epilogue (hi/lo planes out), down GEMM + residual
//   -> final RMSNorm + speech_head GEMV of the LAST position only (t3_hf_backend.py:100-103) -> logits[2][8194]
// so the decode launch that follows finds the slab, the position counter and the pending logits exactly as the
// megakernel's own prefill mode leaves them.
#include "t3_prefill.h"
#include <cstdio>
#include <cstdlib>
#include "s3_gemm.cuh"
#include "t3_rows.cuh"
#include "../../include/cbx.h"

namespace cbx {
namespace {

using namespace t3rows;

// grid (P, B), 1024 threads = (head, d): RoPE (rotate_half pairing d <-> d ^ 32) on q and k in place, q pre-scaled by
// 1/8; K / V rounded to the slab's element type, appended at position pos0 + p and written back so the attention
// below sees exactly what later decode steps will read from the slab.
template <typename KVT>
__global__ void __launch_bounds__(1024) k_rope_kv(float* __restrict__ qkv, int P, int pos0, const float* __restrict__ rope_cos,
                                                  const float* __restrict__ rope_sin, KVT* __restrict__ kc, KVT* __restrict__ vc,
                                                  int layer, int max_pos, long long seq_stride) {
  const int p = blockIdx.x, b = blockIdx.y, head = threadIdx.x >> 6, d = threadIdx.x & 63;
  float* row = qkv + ((size_t)b * P + p) * 3072 + head * 64;
  const float q = row[d], qo = row[d ^ 32], k = row[1024 + d], ko = row[1024 + (d ^ 32)], v = row[2048 + d];
  const float cs = __ldg(rope_cos + (size_t)(pos0 + p) * 32 + (d & 31)), sn = __ldg(rope_sin + (size_t)(pos0 + p) * 32 + (d & 31));
  const float qr = (q * cs + (d < 32 ? -qo : qo) * sn) * 0.125f;
  const float kr = KVS<KVT>::round(k * cs + (d < 32 ? -ko : ko) * sn);
  const float vr = KVS<KVT>::round(v);
  __syncthreads();            // every thread has read its partner before anyone overwrites
  row[d] = qr; row[1024 + d] = kr; row[2048 + d] = vr;
  // batched prefill: row b belongs to utterance b >> 1 (its own slab, seq_stride elements further), CFG row b & 1
  const int bb = seq_stride ? (b & 1) : b;
  const size_t off = (size_t)(seq_stride ? (b >> 1) : 0) * (size_t)seq_stride +
                     (((size_t)(layer * 2 + bb) * NH + head) * max_pos + (pos0 + p)) * HD + d;
  KVS<KVT>::store(kc + off, kr);
  KVS<KVT>::store(vc + off, vr);
}

// Causal attention of the prompt over itself, fp32.  grid (ceil(P / 32), NH, B), 256 threads = 8 warps x 4 queries.
// Keys are walked in tiles of 64 through shared memory (K rows padded to 65 floats); a lane scores keys l and l + 32 for
// the warp's FOUR queries at once (every K / V value read from shared memory feeds four FMAs, the queries come as broadcast
// float4 loads: the first form, one query at a time, was bound by its 1.5 shared-memory loads per FMA — 29 us per launch),
// the warp keeps an online softmax per query, then lane l accumulates output dims l and l + 32.  Output: hi / lo planes.
// Per (query, key) and per (query, dim) the sums run in the same order as before: same bits.
constexpr int QT = 32, KT = 64, KPAD = 65;
constexpr int ATTN_PF_SMEM = (KT * KPAD + KT * 64 + QT * 64 + 8 * KT * 4) * 4;      // 49 408 B: dynamic (just over the static limit)
__global__ void __launch_bounds__(256) k_attn_prefill(const float* __restrict__ qkv, int P, __nv_bfloat16* __restrict__ out,
                                                      long long plane) {
  extern __shared__ __align__(16) float attn_pf_sm[];
  float* Ks = attn_pf_sm;                     // [KT][KPAD]
  float* Vs = Ks + KT * KPAD;                 // [KT][64]   (read along a row: no padding needed)
  float* Qs = Vs + KT * 64;                   // [QT][64]
  float* Ps = Qs + QT * 64;                   // [8 warps][KT][4 queries]
  const int q0 = blockIdx.x * QT, head = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* base = qkv + (size_t)b * P * 3072 + head * 64;
  for (int i = tid; i < QT * 64; i += 256) {
    const int qi = q0 + (i >> 6);
    Qs[i] = qi < P ? base[(size_t)qi * 3072 + (i & 63)] : 0.f;
  }
  float m[4], l[4], o0[4], o1[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { m[j] = -INFINITY; l[j] = 0.f; o0[j] = 0.f; o1[j] = 0.f; }
  const int kmax = min(P, q0 + QT);        // causal: no key beyond the last query of the tile
  const float* qw = Qs + warp * 4 * 64;
  float* pw = Ps + warp * KT * 4;
  for (int k0 = 0; k0 < kmax; k0 += KT) {
    __syncthreads();
    for (int i = tid; i < KT * 64; i += 256) {
      const int kk = k0 + (i >> 6), d = i & 63;
      const bool ok = kk < kmax;
      Ks[(i >> 6) * KPAD + d] = ok ? base[(size_t)kk * 3072 + 1024 + d] : 0.f;
      Vs[(i >> 6) * 64 + d] = ok ? base[(size_t)kk * 3072 + 2048 + d] : 0.f;
    }
    __syncthreads();
    if (q0 + warp * 4 >= P) continue;      // warp-uniform: none of the warp's queries exists (the barriers above were taken)
    float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int d = 0; d < 64; d += 4) {
      float4 q4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q4[j] = *reinterpret_cast<const float4*>(qw + j * 64 + d);
#pragma unroll
      for (int dd = 0; dd < 4; ++dd) {
        const float ka = Ks[lane * KPAD + d + dd], kb = Ks[(lane + 32) * KPAD + d + dd];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float qd = dd == 0 ? q4[j].x : dd == 1 ? q4[j].y : dd == 2 ? q4[j].z : q4[j].w;
          s0[j] += qd * ka;
          s1[j] += qd * kb;
        }
      }
    }
    float corr[4], p0[4], p1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int qi = q0 + warp * 4 + j;      // a query past the prompt: zeros in Qs, its state is never stored
      if (k0 + lane > qi) s0[j] = -INFINITY;
      if (k0 + lane + 32 > qi) s1[j] = -INFINITY;
      float mx = fmaxf(s0[j], s1[j]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const float mn = fmaxf(m[j], mx);    // finite: key k0 <= qi is always present in the first tile
      corr[j] = __expf(m[j] - mn);
      p0[j] = __expf(s0[j] - mn); p1[j] = __expf(s1[j] - mn);
      const float ps = warp_sum(p0[j] + p1[j]);
      l[j] = l[j] * corr[j] + ps;
      m[j] = mn;
    }
    __syncwarp();
    *reinterpret_cast<float4*>(pw + lane * 4) = make_float4(p0[0], p0[1], p0[2], p0[3]);
    *reinterpret_cast<float4*>(pw + (lane + 32) * 4) = make_float4(p1[0], p1[1], p1[2], p1[3]);
    __syncwarp();
    float a0[4], a1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { a0[j] = o0[j] * corr[j]; a1[j] = o1[j] * corr[j]; }
#pragma unroll 8
    for (int kk = 0; kk < KT; ++kk) {
      const float4 p4 = *reinterpret_cast<const float4*>(pw + kk * 4);
      const float va = Vs[kk * 64 + lane], vb = Vs[kk * 64 + lane + 32];
      a0[0] += p4.x * va; a1[0] += p4.x * vb;
      a0[1] += p4.y * va; a1[1] += p4.y * vb;
      a0[2] += p4.z * va; a1[2] += p4.z * vb;
      a0[3] += p4.w * va; a1[3] += p4.w * vb;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { o0[j] = a0[j]; o1[j] = a1[j]; }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int qi = q0 + warp * 4 + j;
    if (qi >= P) continue;
    const float inv = 1.0f / l[j];
    const float v0 = o0[j] * inv, v1 = o1[j] * inv;
    const __nv_bfloat16 h0 = __float2bfloat16_rn(v0), h1 = __float2bfloat16_rn(v1);
    __nv_bfloat16* d = out + ((size_t)b * P + qi) * H + head * 64;
    d[lane] = h0; d[lane + 32] = h1;
    d[plane + lane] = __float2bfloat16_rn(v0 - __bfloat162float(h0));
    d[plane + lane + 32] = __float2bfloat16_rn(v1 - __bfloat162float(h1));
  }
}

// final RMSNorm of the last position of each CFG row + speech_head GEMV: one warp per vocabulary row
__global__ void __launch_bounds__(256) k_head(const float* __restrict__ x, int P, int B, const float* __restrict__ gain,
                                              const __nv_bfloat16* __restrict__ w, float* __restrict__ logits, int ld) {
  __shared__ float xs[2][H];
  __shared__ float red[2][8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  x += (size_t)blockIdx.y * 2 * P * H;          // batched prefill: blockIdx.y = utterance (its CFG pair of rows / logits rows)
  logits += (size_t)blockIdx.y * 2 * ld;
  for (int b = 0; b < B; ++b) {
    const float4 v = *reinterpret_cast<const float4*>(x + ((size_t)b * P + (P - 1)) * H + tid * 4);
    float s = warp_sum(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
    if (lane == 0) red[b][warp] = s;
    *reinterpret_cast<float4*>(&xs[b][tid * 4]) = v;
  }
  __syncthreads();
  for (int b = 0; b < B; ++b) {
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[b][i];
    const float sc = rsqrtf(tot * (1.0f / H) + RMS_EPS);
    const float4 g = __ldg(reinterpret_cast<const float4*>(gain + tid * 4));
    float4 v = *reinterpret_cast<float4*>(&xs[b][tid * 4]);
    v.x = g.x * (v.x * sc); v.y = g.y * (v.y * sc); v.z = g.z * (v.z * sc); v.w = g.w * (v.w * sc);
    *reinterpret_cast<float4*>(&xs[b][tid * 4]) = v;
  }
  __syncthreads();
  const int row = blockIdx.x * 8 + warp;
  if (row >= VOCAB) return;
  const __nv_bfloat16* wr = w + (size_t)row * H;
  float a0 = 0.f, a1 = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = i * 256 + lane * 8;
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(wr + k));
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h2[e]);
      a0 += f.x * xs[0][k + 2 * e] + f.y * xs[0][k + 2 * e + 1];
      if (B == 2) a1 += f.x * xs[1][k + 2 * e] + f.y * xs[1][k + 2 * e + 1];
    }
  }
  a0 = warp_sum(a0); a1 = warp_sum(a1);
  if (lane == 0) {
    logits[row] = a0;
    if (B == 2) logits[ld + row] = a1;
  }
}

__global__ void k_set_pos(int* st, int pos) { st[0] = pos; st[1] = 0; st[2] = 0; }

// Alignment analyzer rows of the prompt (alignment_stream_analyzer.py:73-75,96-101): softmax(q k^T) of CFG row 0 at query
// position P - rows0 + blockIdx.x, averaged over the 16 heads, for the text columns.  qkv holds q (RoPE, pre-scaled by
// 1/8) and k (RoPE, rounded like the slab) of this layer.  512 threads = 16 warps = heads for the softmax statistics.
__global__ void __launch_bounds__(512) k_align_rows(const float* __restrict__ qkv, int P, int rows0, int al_i, int al_n,
                                                    float* __restrict__ out, int cap) {
  __shared__ float qs[NH][HD];
  __shared__ float ml[NH][2];
  const int r = blockIdx.x, qi = P - rows0 + r, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  qs[warp][lane] = qkv[(size_t)qi * 3072 + warp * HD + lane];
  qs[warp][lane + 32] = qkv[(size_t)qi * 3072 + warp * HD + lane + 32];
  __syncthreads();
  auto score = [&](int h, int k) {
    const float* kr = qkv + (size_t)k * 3072 + 1024 + h * HD;
    float s = 0.f;
#pragma unroll 16
    for (int d = 0; d < HD; ++d) s += qs[h][d] * kr[d];
    return s;
  };
  float m = -INFINITY;
  for (int k = lane; k <= qi; k += 32) m = fmaxf(m, score(warp, k));
  m = warp_max(m);
  float l = 0.f;
  for (int k = lane; k <= qi; k += 32) l += __expf(score(warp, k) - m);
  l = warp_sum(l);
  if (lane == 0) { ml[warp][0] = m; ml[warp][1] = 1.0f / (l * (float)NH); }
  __syncthreads();
  for (int c = tid; c < al_n; c += 512) {
    const int k = al_i + c;
    float acc = 0.f;
    if (k <= qi)
      for (int h = 0; h < NH; ++h) acc += __expf(score(h, k) - ml[h][0]) * ml[h][1];
    out[(size_t)r * cap + c] = acc;
  }
}

}  // namespace

size_t t3_prefill_workspace_bytes_rows(size_t rows) {
  return rows * H * 4 + rows * 3072 * 4 + 2 * rows * H * 2 + 2 * rows * INTER * 2 + 1024;
}
size_t t3_prefill_workspace_bytes(int max_pos) { return t3_prefill_workspace_bytes_rows(2 * (size_t)max_pos); }

int t3_prefill_gemm(const T3PrefillArgs& a, cudaStream_t st) {
  static CbxPerDeviceOnce attn_once;
  if (attn_once.first()) CBX_CUDA(cudaFuncSetAttribute(k_attn_prefill, cudaFuncAttributeMaxDynamicSharedMemorySize, ATTN_PF_SMEM));
  const int rows = a.B * a.P;
  const size_t maxrows = a.ws_rows > 0 ? (size_t)a.ws_rows : 2 * (size_t)a.max_pos;
  const bool many = a.nseq > 1;
  CBX_CHECK((size_t)rows <= maxrows, "t3_prefill_gemm: %d rows exceed the workspace (%zu rows)", rows, maxrows);
  CBX_CHECK(!many || (a.B == 2 * a.nseq && a.kv_seq_stride > 0 && a.al_layer < 0), "t3_prefill_gemm: inconsistent batched arguments");
  const long long kvs = many ? a.kv_seq_stride : 0;
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  float* x = reinterpret_cast<float*>(ws);
  float* qkv = reinterpret_cast<float*>(ws + maxrows * H * 4);
  __nv_bfloat16* hl = reinterpret_cast<__nv_bfloat16*>(ws + maxrows * H * 4 + maxrows * 3072 * 4);
  __nv_bfloat16* act = hl + 2 * maxrows * H;
  const long long hl_plane = (long long)rows * H, act_plane = (long long)rows * INTER;
  CBX_CUDA(cudaMemcpyAsync(x, a.embeds, (size_t)rows * H * 4, cudaMemcpyDeviceToDevice, st));
  const __nv_bfloat16* wp = static_cast<const __nv_bfloat16*>(a.wplain);
  const size_t per_layer = (size_t)3072 * H + (size_t)H * H + (size_t)8192 * H + (size_t)H * INTER;
  // column tile / split-K cluster of the q|k|v, o_proj and down GEMMs: the prompt is 2 row tiles, so narrow tiles put ~100 CTAs
  // on the SMs and every CTA streams a smaller weight panel (the widest tile left 24-48 CTAs ingesting 0.5 MB each).
  // CBX_T3P_TILES="qkv_t,qkv_s,o_t,o_s,dn_t,dn_s" overrides; 0 = the GEMM's own choice.
  struct PTiles { int qkv_t, qkv_s, o_t, o_s, dn_t, dn_s; };
  static const PTiles pt = [] {
    PTiles c{128, 2, 64, 4, 64, 4};      // measured at P = 102: 2.35 ms per prefill (widest tiles: 2.91; 64,2,32,4,32,4: 3.23)
    if (const char* v = getenv("CBX_T3P_TILES")) {
      int t[6];
      if (sscanf(v, "%d,%d,%d,%d,%d,%d", t, t + 1, t + 2, t + 3, t + 4, t + 5) == 6) c = PTiles{t[0], t[1], t[2], t[3], t[4], t[5]};
    }
    return c;
  }();
  // many utterances: thousands of rows fill the GPU with the widest tiles, no split-K
  const PTiles ptl = rows > 1024 ? PTiles{0, 1, 0, 1, 0, 1} : pt;
  for (int l = 0; l < a.n_layers; ++l) {
    const __nv_bfloat16* wqkv = wp + l * per_layer;
    const __nv_bfloat16* wo = wqkv + (size_t)3072 * H;
    const __nv_bfloat16* wgu = wo + (size_t)H * H;
    const __nv_bfloat16* wdn = wgu + (size_t)8192 * H;
    GemmDesc d;
    GemmEpi e;
    k_rms_hilo<<<rows, 256, 0, st>>>(x, a.ln1 + (size_t)l * H, hl, hl_plane);
    e = GemmEpi{}; e.out_f32 = qkv; e.ld_f32 = 3072;
    if (gemm_make_hilo_ex(&d, hl, H, hl_plane, wqkv, rows, 3072, 3072, H, e, ptl.qkv_t, ptl.qkv_s) || gemm_launch(d, st)) return CBX_ERR;
    if (a.kv_fp32)
      k_rope_kv<float><<<dim3(a.P, a.B), 1024, 0, st>>>(qkv, a.P, 0, a.rope_cos, a.rope_sin, static_cast<float*>(a.kcache),
                                                        static_cast<float*>(a.vcache), l, a.max_pos, kvs);
    else
      k_rope_kv<__nv_bfloat16><<<dim3(a.P, a.B), 1024, 0, st>>>(qkv, a.P, 0, a.rope_cos, a.rope_sin,
                                                               static_cast<__nv_bfloat16*>(a.kcache),
                                                               static_cast<__nv_bfloat16*>(a.vcache), l, a.max_pos, kvs);
    if (l == a.al_layer && a.al_out)
      k_align_rows<<<a.al_rows0, 512, 0, st>>>(qkv, a.P, a.al_rows0, a.al_i, a.al_n, a.al_out, a.al_cap);
    k_attn_prefill<<<dim3((a.P + QT - 1) / QT, NH, a.B), 256, ATTN_PF_SMEM, st>>>(qkv, a.P, hl, hl_plane);
    e = GemmEpi{}; e.resid = x; e.ld_resid = H; e.out_f32 = x; e.ld_f32 = H;
    if (gemm_make_hilo_ex(&d, hl, H, hl_plane, wo, rows, H, H, H, e, ptl.o_t, ptl.o_s) || gemm_launch(d, st)) return CBX_ERR;
    k_rms_hilo<<<rows, 256, 0, st>>>(x, a.ln2 + (size_t)l * H, hl, hl_plane);
    e = GemmEpi{}; e.swiglu = 1; e.out_hilo = act; e.ld_hilo = INTER; e.hilo_plane = act_plane;
    if (gemm_make_hilo(&d, hl, H, hl_plane, wgu, rows, 8192, H, e) || gemm_launch(d, st)) return CBX_ERR;
    e = GemmEpi{}; e.resid = x; e.ld_resid = H; e.out_f32 = x; e.ld_f32 = H;
    if (gemm_make_hilo_ex(&d, act, INTER, act_plane, wdn, rows, H, H, INTER, e, ptl.dn_t, ptl.dn_s) || gemm_launch(d, st)) return CBX_ERR;
  }
  k_head<<<dim3((VOCAB + 7) / 8, many ? a.nseq : 1), 256, 0, st>>>(x, a.P, many ? 2 : a.B, a.final_norm, wp + a.n_layers * per_layer,
                                                                     a.logits, a.logits_ld > 0 ? a.logits_ld : VOCAB);
  k_set_pos<<<1, 1, 0, st>>>(a.state, a.P);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

}  // namespace cbx
