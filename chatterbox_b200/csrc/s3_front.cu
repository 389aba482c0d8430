// S3Gen flow front-end for sm_100a: the part of CausalMaskedDiffWithXvec.inference (reference
// src/chatterbox/models/s3gen/flow.py:236-282) that runs once per utterance before the flow-matching loop —
// x-vector projection, token embedding, UpsampleConformerEncoder.forward (transformer/upsample_encoder.py:237-304:
// LinearNoSubsampling + ESPnet relative positional encoding, PreLookaheadLayer, 6 conformer blocks at 25 Hz, nearest x2
// upsample + conv, 4 blocks at 50 Hz, final LayerNorm), encoder_proj, and the prompt-mel `cond` tensor.
//
//   * every Linear and convolution runs on the tcgen05 GEMM (s3_gemm.cu) in hi/lo mode: activations as two bf16 planes
//     (~16 mantissa bits) against bf16 weights, fp32 accumulation in TMEM; convolutions through an im2col of the (tiny)
//     activation that writes the planes directly;
//   * relative-position attention (transformer/attention.py:249-330) is one kernel: scores (q + u) k^T + (q + v) p^T with
//     `rel_shift` (attention.py:225-247) applied as an INDEX transform, bd[i][j] = ((q_i + v) . p[T - 1 - i + j]) — the
//     reference materialises the (1, 8, T, 2T - 1) score tensor and re-views it; here a (queries x keys) tile needs the 63
//     consecutive rows of p that its diagonals touch — online softmax, fp32 on CUDA cores (8 heads x 64, T <= ~2000);
//   * `linear_pos(pos_emb)` does not depend on the input: it is computed once per (layer, T) when the plan for a length is
//     built and cached with it.
#include "s3_gemm.cuh"
#include "../../include/cbx.h"
#include <math.h>
#include <map>
#include <vector>

namespace {

using namespace cbx;

constexpr int D = 512, NH = 8, DK = 64, FF = 2048, NMEL = 80, XV = 192;
constexpr int N_LAYERS0 = 6, N_LAYERS1 = 4, N_LAYERS = N_LAYERS0 + N_LAYERS1;

__device__ __forceinline__ void store_hilo4(__nv_bfloat16* dst, long long plane, float a, float b, float c, float d) {
  uint32_t h0, l0, h1, l1;
  split_hi_lo(a, b, h0, l0);
  split_hi_lo(c, d, h1, l1);
  *reinterpret_cast<uint2*>(dst) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(dst + plane) = make_uint2(l0, l1);
}

// x[t] = embedding[clamp(token[t], 0)] as hi / lo planes [2][T][512]      (flow.py:253-255)
__global__ void __launch_bounds__(128) k_sf_embed(const long long* __restrict__ prompt, int Tp, const long long* __restrict__ tok, int T,
                                                  const float* __restrict__ emb, __nv_bfloat16* __restrict__ out, long long plane) {
  const int t = blockIdx.x, c = threadIdx.x;
  long long id = t < Tp ? prompt[t] : tok[t - Tp];
  if (id < 0) id = 0;
  const float4 v = __ldg(reinterpret_cast<const float4*>(emb + (size_t)id * D) + c);
  store_hilo4(out + (size_t)t * D + c * 4, plane, v.x, v.y, v.z, v.w);
}

// LayerNorm over 512 channels, one row per CTA (128 threads x 4).  out_f32 (nullable) = scale * LN(x); out_hilo (nullable) =
// the same as hi / lo planes.  eps: 1e-5 (embed / after_norm) or 1e-12 (conformer blocks, encoder_layer.py).
__global__ void __launch_bounds__(128) k_sf_ln(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b,
                                               float eps, float scale, float* __restrict__ out_f32,
                                               __nv_bfloat16* __restrict__ out_hilo, long long plane) {
  __shared__ float red[2][4];
  const int r = blockIdx.x, t = threadIdx.x;
  const float4 v = reinterpret_cast<const float4*>(x + (size_t)r * D)[t];
  float s = warp_sum(v.x + v.y + v.z + v.w);
  if ((t & 31) == 0) red[0][t >> 5] = s;
  __syncthreads();
  const float mean = (red[0][0] + red[0][1] + red[0][2] + red[0][3]) * (1.0f / D);
  const float dx = v.x - mean, dy = v.y - mean, dz = v.z - mean, dw = v.w - mean;
  float q = warp_sum(dx * dx + dy * dy + dz * dz + dw * dw);
  if ((t & 31) == 0) red[1][t >> 5] = q;
  __syncthreads();
  const float rstd = rsqrtf((red[1][0] + red[1][1] + red[1][2] + red[1][3]) * (1.0f / D) + eps);
  const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + t), bb = __ldg(reinterpret_cast<const float4*>(b) + t);
  const float y0 = (dx * rstd * gg.x + bb.x) * scale, y1 = (dy * rstd * gg.y + bb.y) * scale;
  const float y2 = (dz * rstd * gg.z + bb.z) * scale, y3 = (dw * rstd * gg.w + bb.w) * scale;
  if (out_f32) reinterpret_cast<float4*>(out_f32 + (size_t)r * D)[t] = make_float4(y0, y1, y2, y3);
  if (out_hilo) store_hilo4(out_hilo + (size_t)r * D + t * 4, plane, y0, y1, y2, y3);
}

// act(x) as hi / lo planes: act 0 = identity, 1 = SiLU (feed-forward, positionwise_feed_forward.py), 2 = leaky_relu(0.01)
__global__ void __launch_bounds__(256) k_sf_act_hilo(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, long long plane,
                                                     long long n4, int act) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= n4) return;
  float4 v = reinterpret_cast<const float4*>(x)[i];
  if (act == 1) {
    v.x = v.x / (1.f + __expf(-v.x)); v.y = v.y / (1.f + __expf(-v.y)); v.z = v.z / (1.f + __expf(-v.z)); v.w = v.w / (1.f + __expf(-v.w));
  } else if (act == 2) {
    v.x = v.x > 0.f ? v.x : 0.01f * v.x; v.y = v.y > 0.f ? v.y : 0.01f * v.y; v.z = v.z > 0.f ? v.z : 0.01f * v.z; v.w = v.w > 0.f ? v.w : 0.01f * v.w;
  }
  store_hilo4(out + i * 4, plane, v.x, v.y, v.z, v.w);
}

// im2col of a [Tin][512] fp32 activation for a k-tap convolution -> hi / lo planes [2][Tout][taps * 512], column = tap * 512 + c:
// source row of output row t and tap j = (t + j - left) / up (nearest-neighbour upsampling by `up` folded in); rows outside
// [0, Tin * up) are zero padding.  act as above (applied to the source value).
__global__ void __launch_bounds__(128) k_sf_im2col(const float* __restrict__ x, int Tin, int up, int taps, int left, int act,
                                                   __nv_bfloat16* __restrict__ out, long long plane) {
  const int t = blockIdx.x, j = blockIdx.y, c = threadIdx.x;
  const int su = t + j - left;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (su >= 0 && su < Tin * up) {
    v = reinterpret_cast<const float4*>(x + (size_t)(su / up) * D)[c];
    if (act == 2) { v.x = v.x > 0.f ? v.x : 0.01f * v.x; v.y = v.y > 0.f ? v.y : 0.01f * v.y; v.z = v.z > 0.f ? v.z : 0.01f * v.z; v.w = v.w > 0.f ? v.w : 0.01f * v.w; }
  }
  store_hilo4(out + ((size_t)t * taps + j) * D + c * 4, plane, v.x, v.y, v.z, v.w);
}

// EspnetRelPositionalEncoding.position_encoding (embedding.py:224-294) for length T as hi / lo planes [2][2T-1][512]:
// row m is relative position T - 1 - m; even channels sin(pos * div), odd channels cos(pos * div)
__global__ void __launch_bounds__(256) k_sf_posemb(int T, __nv_bfloat16* __restrict__ out, long long plane) {
  const int m = blockIdx.x, i = threadIdx.x;          // i: channel pair
  const float pos = (float)(T - 1 - m);
  const float div = expf((float)(2 * i) * -(logf(10000.0f) / (float)D));
  const float a = pos * div;
  const float s = sinf(a), c = cosf(a);
  const __nv_bfloat16 hs = __float2bfloat16_rn(s), hc = __float2bfloat16_rn(c);
  __nv_bfloat16* d = out + (size_t)m * D + 2 * i;
  d[0] = hs; d[1] = hc;
  d[plane] = __float2bfloat16_rn(s - __bfloat162float(hs));
  d[plane + 1] = __float2bfloat16_rn(c - __bfloat162float(hc));
}

// ---- relative-position multi-head attention (transformer/attention.py:249-330), 8 heads x 64, one un-padded utterance ----
// CTA = (32 queries, head).  qkv [T][1536] fp32 (q | k | v with bias), pp [2T-1][512] fp32 = linear_pos(pos_emb), u / v biases
// [8][64].  Key tiles of 32: K, V tiles and the 63 rows of p the tile's diagonals touch go through shared memory; scores
// 4 per thread, online softmax per query row (8 lanes share a row), output 8 dims per thread.  Output: hi / lo planes [T][512].
constexpr int RA_Q = 32, RA_K = 32, RA_LD = 65;
__global__ void __launch_bounds__(256) k_sf_relattn(const float* __restrict__ qkv, const float* __restrict__ pp, const float* __restrict__ bu,
                                                    const float* __restrict__ bv, int T, __nv_bfloat16* __restrict__ out, long long plane) {
  // 45.6 KB of static shared memory: one Q tile (the two position biases are added on the fly)
  __shared__ float Qs[RA_Q * RA_LD], Ks[RA_K * RA_LD], Vs[RA_K * RA_LD], Ps[(RA_Q + RA_K - 1) * RA_LD], Ss[RA_Q * (RA_K + 1)];
  __shared__ float bus[DK], bvs[DK];
  const int i0 = blockIdx.x * RA_Q, h = blockIdx.y, tid = threadIdx.x;
  for (int e = tid; e < RA_Q * DK; e += 256) {
    const int r = e >> 6, d = e & 63;
    Qs[r * RA_LD + d] = (i0 + r < T) ? qkv[(size_t)(i0 + r) * 1536 + h * DK + d] : 0.f;
  }
  if (tid < DK) { bus[tid] = __ldg(bu + h * DK + tid); bvs[tid] = __ldg(bv + h * DK + tid); }
  const int qi = tid >> 3, sub = tid & 7;            // query row of this thread; keys sub*4..+3 / output dims sub*8..+7
  float m = -INFINITY, l = 0.f, o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = 0.f;
  for (int j0 = 0; j0 < T; j0 += RA_K) {
    __syncthreads();                                 // previous tile's readers are done (and Qs / the biases are written)
    for (int e = tid; e < RA_K * DK; e += 256) {
      const int r = e >> 6, d = e & 63;
      const bool ok = j0 + r < T;
      Ks[r * RA_LD + d] = ok ? qkv[(size_t)(j0 + r) * 1536 + 512 + h * DK + d] : 0.f;
      Vs[r * RA_LD + d] = ok ? qkv[(size_t)(j0 + r) * 1536 + 1024 + h * DK + d] : 0.f;
    }
    // p rows m = T - 1 - i + j for i in [i0, i0 + 32), j in [j0, j0 + 32): local row (j - j0) - (i - i0) + 31  <->  m0 + local
    const int m0 = T - 1 - (i0 + RA_Q - 1) + j0;
    for (int e = tid; e < (RA_Q + RA_K - 1) * DK; e += 256) {
      const int r = e >> 6, d = e & 63, mm = m0 + r;
      Ps[r * RA_LD + d] = (mm >= 0 && mm < 2 * T - 1) ? pp[(size_t)mm * D + h * DK + d] : 0.f;
    }
    __syncthreads();
    float s[4] = {0.f, 0.f, 0.f, 0.f};
    const float* qr = Qs + qi * RA_LD;
#pragma unroll 8
    for (int d = 0; d < DK; ++d) {
      const float q = qr[d], a = q + bus[d], b = q + bvs[d];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int j = sub * 4 + jj;
        s[jj] += a * Ks[j * RA_LD + d] + b * Ps[(j - qi + RA_Q - 1) * RA_LD + d];
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      s[jj] = (j0 + sub * 4 + jj < T) ? s[jj] * 0.125f : -INFINITY;      // 1 / sqrt(d_k); keys past T masked
      mx = fmaxf(mx, s[jj]);
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
    const float mn = fmaxf(m, mx);                   // finite: the first key of every tile exists
    const float corr = __expf(m - mn);
    float ls = 0.f;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const float pe = __expf(s[jj] - mn);
      Ss[qi * (RA_K + 1) + sub * 4 + jj] = pe;
      ls += pe;
    }
    ls += __shfl_xor_sync(0xffffffffu, ls, 1);
    ls += __shfl_xor_sync(0xffffffffu, ls, 2);
    ls += __shfl_xor_sync(0xffffffffu, ls, 4);
    l = l * corr + ls;
    m = mn;
    __syncwarp();                                    // the 8 lanes of a row live in one warp: its Ss row is complete
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] *= corr;
#pragma unroll 4
    for (int j = 0; j < RA_K; ++j) {
      const float pj = Ss[qi * (RA_K + 1) + j];
      const float* vr = Vs + j * RA_LD + sub * 8;
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] += pj * vr[e];
    }
  }
  if (i0 + qi < T) {
    const float inv = 1.0f / l;
    __nv_bfloat16* dst = out + (size_t)(i0 + qi) * D + h * DK + sub * 8;
    store_hilo4(dst, plane, o[0] * inv, o[1] * inv, o[2] * inv, o[3] * inv);
    store_hilo4(dst + 4, plane, o[4] * inv, o[5] * inv, o[6] * inv, o[7] * inv);
  }
}

// ---- the same attention on the tensor cores: mma.sync m16n8k8 TF32 (fp32 operands straight from shared memory, 10-bit mantissa
// products, fp32 accumulation).  CTA = (64 queries, head), 4 warps x 16 query rows; key tiles of 64, double-buffered with
// cp.async: K, V and the 127 rows of p the tile's diagonals touch.  Per warp and tile: content scores (q + u) K^T, 16 x 64;
// position scores (q + v) P_band^T, 16 x 80 (the band columns its rows can reach), written to a per-warp scratch and read back
// along the diagonals (rel_shift is the index m = T - 1 - i + j, i.e. band column j - i + const); online softmax in the
// accumulator layout; P V with the probabilities re-used as the A operand in place (the two key columns a thread holds are
// presented as A columns t and t + 4, and the V rows are fetched in the same order).  (q + u), (q + v) live in registers as A
// fragments for the whole key loop.  CBX_RELATTN_TC=0 selects the CUDA-core kernel above. ----
constexpr int RT_Q = 64, RT_K = 64, RT_LD = 68, RT_BAND = 128, RT_BLD = 84;
constexpr int RT_STAGE = (2 * RT_K + RT_BAND) * RT_LD;              // floats: K, V, band
constexpr int RT_SMEM = (2 * RT_STAGE + 4 * 16 * RT_BLD) * 4;
__device__ __forceinline__ void rt_cp16(float* dst, const float* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ uint32_t rt_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void rt_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__global__ void __launch_bounds__(128) k_sf_relattn_tc(const float* __restrict__ qkv, const float* __restrict__ pp, const float* __restrict__ bu,
                                                       const float* __restrict__ bv, int T, __nv_bfloat16* __restrict__ out, long long plane) {
  extern __shared__ __align__(16) float rt_smem[];
  float* Bs = rt_smem + 2 * RT_STAGE;
  const int i0 = blockIdx.x * RT_Q, h = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  auto load_tile = [&](int kt, int buf) {
    float* Ks = rt_smem + buf * RT_STAGE;
    float* Vs = Ks + RT_K * RT_LD;
    float* Ps = Vs + RT_K * RT_LD;
    const int j0 = kt * RT_K;
    for (int e = tid; e < RT_K * 16; e += 128) {
      const int r = e >> 4, c = (e & 15) * 4;
      const bool ok = j0 + r < T;
      const float* src = qkv + (size_t)(ok ? j0 + r : 0) * 1536 + h * DK + c;
      rt_cp16(Ks + r * RT_LD + c, src + 512, ok);
      rt_cp16(Vs + r * RT_LD + c, src + 1024, ok);
    }
    const int m0 = T - 1 - (i0 + RT_Q - 1) + j0;       // band row bl <-> p row m0 + bl; bl = (j - j0) - (i - i0) + 63
    for (int e = tid; e < RT_BAND * 16; e += 128) {
      const int r = e >> 4, c = (e & 15) * 4, mm = m0 + r;
      const bool ok = mm >= 0 && mm < 2 * T - 1;
      rt_cp16(Ps + r * RT_LD + c, pp + (size_t)(ok ? mm : 0) * D + h * DK + c, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int ntiles = (T + RT_K - 1) / RT_K;
  load_tile(0, 0);
  // (q + u), (q + v) of this warp's 16 rows as A fragments: a0 = (g, t), a1 = (g + 8, t), a2 = (g, t + 4), a3 = (g + 8, t + 4)
  uint32_t qu[8][4], qv[8][4];
  {
    const int r0 = i0 + warp * 16 + g, r1 = r0 + 8;
    const float* q0 = qkv + (size_t)(r0 < T ? r0 : 0) * 1536 + h * DK;
    const float* q1 = qkv + (size_t)(r1 < T ? r1 : 0) * 1536 + h * DK;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int c0 = ks * 8 + t, c1 = c0 + 4;
      const float u0 = __ldg(bu + h * DK + c0), u1 = __ldg(bu + h * DK + c1), v0 = __ldg(bv + h * DK + c0), v1 = __ldg(bv + h * DK + c1);
      const float a = r0 < T ? q0[c0] : 0.f, b2 = r1 < T ? q1[c0] : 0.f, c = r0 < T ? q0[c1] : 0.f, d2 = r1 < T ? q1[c1] : 0.f;
      qu[ks][0] = rt_tf32(a + u0); qu[ks][1] = rt_tf32(b2 + u0); qu[ks][2] = rt_tf32(c + u1); qu[ks][3] = rt_tf32(d2 + u1);
      qv[ks][0] = rt_tf32(a + v0); qv[ks][1] = rt_tf32(b2 + v0); qv[ks][2] = rt_tf32(c + v1); qv[ks][3] = rt_tf32(d2 + v1);
    }
  }
  float o[8][4], m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float* Bw = Bs + warp * 16 * RT_BLD;
  const int b0 = 48 - 16 * warp;                     // first band column this warp's rows can reach
  for (int kt = 0; kt < ntiles; ++kt) {
    if (kt + 1 < ntiles) load_tile(kt + 1, (kt + 1) & 1);
    else asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    const float* Ks = rt_smem + (kt & 1) * RT_STAGE;
    const float* Vs = Ks + RT_K * RT_LD;
    const float* Ps = Vs + RT_K * RT_LD;
    // position scores of the band: 16 rows x 80 columns -> per-warp scratch
#pragma unroll
    for (int nt = 0; nt < 10; ++nt) {
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      const float* pr = Ps + (b0 + nt * 8 + g) * RT_LD + t;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) rt_mma(c, qv[ks], __float_as_uint(pr[ks * 8]), __float_as_uint(pr[ks * 8 + 4]));
      *reinterpret_cast<float2*>(Bw + g * RT_BLD + nt * 8 + 2 * t) = make_float2(c[0], c[1]);
      *reinterpret_cast<float2*>(Bw + (g + 8) * RT_BLD + nt * 8 + 2 * t) = make_float2(c[2], c[3]);
    }
    __syncwarp();
    // content scores + the band along the diagonals: row g takes scratch column (key - g + 15), row g + 8 column (key - g + 7)
    float s[8][4];
    const int j0 = kt * RT_K;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      const float* kr = Ks + (nt * 8 + g) * RT_LD + t;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) rt_mma(c, qu[ks], __float_as_uint(kr[ks * 8]), __float_as_uint(kr[ks * 8 + 4]));
      const int key = nt * 8 + 2 * t;
      const float* d0 = Bw + g * RT_BLD + key - g + 15;
      const float* d1 = Bw + (g + 8) * RT_BLD + key - g + 7;
      s[nt][0] = (j0 + key < T) ? (c[0] + d0[0]) * 0.125f : -INFINITY;
      s[nt][1] = (j0 + key + 1 < T) ? (c[1] + d0[1]) * 0.125f : -INFINITY;
      s[nt][2] = (j0 + key < T) ? (c[2] + d1[0]) * 0.125f : -INFINITY;
      s[nt][3] = (j0 + key + 1 < T) ? (c[3] + d1[1]) * 0.125f : -INFINITY;
    }
    // online softmax (rows g and g + 8; the four lanes of a quad share a row)
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) { mx[0] = fmaxf(mx[0], fmaxf(s[nt][0], s[nt][1])); mx[1] = fmaxf(mx[1], fmaxf(s[nt][2], s[nt][3])); }
    float corr[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float mn = fmaxf(m_run[r], mx[r]);       // finite: the first key of every tile exists
      corr[r] = __expf(m_run[r] - mn);
      m_run[r] = mn;
    }
    float ls[2] = {0.f, 0.f};
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = __expf(s[nt][0] - m_run[0]); s[nt][1] = __expf(s[nt][1] - m_run[0]);
      s[nt][2] = __expf(s[nt][2] - m_run[1]); s[nt][3] = __expf(s[nt][3] - m_run[1]);
      ls[0] += s[nt][0] + s[nt][1]; ls[1] += s[nt][2] + s[nt][3];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) l_run[r] = l_run[r] * corr[r] + ls[r];
#pragma unroll
    for (int i = 0; i < 8; ++i) { o[i][0] *= corr[0]; o[i][1] *= corr[0]; o[i][2] *= corr[1]; o[i][3] *= corr[1]; }
    // O += P V: k-step = key block nt; A columns t / t + 4 <-> keys 8 nt + 2 t / + 1, V rows fetched in the same order
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const uint32_t a[4] = {rt_tf32(s[nt][0]), rt_tf32(s[nt][2]), rt_tf32(s[nt][1]), rt_tf32(s[nt][3])};
      const float* vr = Vs + (nt * 8 + 2 * t) * RT_LD + g;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) rt_mma(o[nd], a, __float_as_uint(vr[nd * 8]), __float_as_uint(vr[RT_LD + nd * 8]));
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = i0 + warp * 16 + g + r * 8;
    if (row < T) {
      const float inv = 1.0f / l_run[r];
      __nv_bfloat16* dst = out + (size_t)row * D + h * DK + 2 * t;
#pragma unroll
      for (int nd = 0; nd < 8; ++nd) {
        uint32_t hi, lo;
        split_hi_lo(o[nd][r * 2] * inv, o[nd][r * 2 + 1] * inv, hi, lo);
        *reinterpret_cast<uint32_t*>(dst + nd * 8) = hi;
        *reinterpret_cast<uint32_t*>(dst + nd * 8 + plane) = lo;
      }
    }
  }
}

// spks = Linear(192 -> 80)(F.normalize(embedding))     (flow.py:250-251)
__global__ void __launch_bounds__(96) k_sf_spk(const float* __restrict__ emb, const float* __restrict__ w, const float* __restrict__ b,
                                               float* __restrict__ out) {
  __shared__ float e[XV];
  __shared__ float nrm;
  const int t = threadIdx.x;
  e[t] = emb[t]; e[t + 96] = emb[t + 96];
  __syncthreads();
  if (t == 0) {
    float s = 0.f;
    for (int i = 0; i < XV; ++i) s += e[i] * e[i];
    nrm = fmaxf(sqrtf(s), 1e-12f);                   // F.normalize: x / max(||x||, eps)
  }
  __syncthreads();
  if (t < NMEL) {
    float a = 0.f;
    for (int i = 0; i < XV; ++i) a += __ldg(w + (size_t)t * XV + i) * (e[i] / nrm);
    out[t] = a + __ldg(b + t);
  }
}

// mu [80][L] <- h [L][80] (channel-first for the CFM entry points); cond [80][L] = prompt mel in the first Lp frames, zeros after
__global__ void __launch_bounds__(NMEL) k_sf_finish(const float* __restrict__ hproj, const float* __restrict__ prompt_feat, int Lp, int L,
                                                    float* __restrict__ mu, float* __restrict__ cond) {
  const int t = blockIdx.x, c = threadIdx.x;
  mu[(size_t)c * L + t] = hproj[(size_t)t * NMEL + c];
  cond[(size_t)c * L + t] = t < Lp ? prompt_feat[(size_t)t * NMEL + c] : 0.f;
}

}  // namespace

// =================================================================================================
// host side of the C-ABI (include/cbx.h, section 2b)
// =================================================================================================
namespace {
// bf16 matrices [N][K] (pack order of chatterbox_b200/flow_frontend.py:pack_front)
enum { BF_EMBED = 0, BF_PL1 = 1, BF_PL2 = 2, BF_LAYER = 3 /* + 5 i: qkv, pos, out, w1, w2 */, BF_UPCONV = BF_LAYER + 5 * N_LAYERS,
       BF_UPEMBED, BF_PROJ, BF_COUNT };
// fp32 vectors / tables
enum { F_TABLE = 0, F_EMBED_B, F_EMBED_G, F_EMBED_LB, F_PL1_B, F_PL2_B,
       F_LAYER /* + 10 i: g1, b1, bqkv, bo, pos_u, pos_v, g2, b2, bw1, bw2 */, F_UPCONV_B = F_LAYER + 10 * N_LAYERS, F_UPEMBED_B, F_UPEMBED_G,
       F_UPEMBED_LB, F_AFTER_G, F_AFTER_B, F_PROJ_B, F_SPK_W, F_SPK_B, F_COUNT };
constexpr int FE_MAX_PLANS = 8;

struct FePlan {
  int T0;                      // tokens (prompt + utterance); the second half of the encoder runs at 2 T0
  float* pp;                   // linear_pos(pos_emb) of every layer: layers 0-5 [2 T0 - 1][512] each, then layers 6-9 [4 T0 - 1][512]
  GemmDesc embed, pl1, pl2, upconv, upembed, proj;
  GemmDesc qkv[N_LAYERS], out[N_LAYERS], w1[N_LAYERS], w2[N_LAYERS];
  long long stamp;
};
}  // namespace

struct cbx_flowfe {
  int device, max_tokens;
  const __nv_bfloat16* wbf; const float* wf;
  std::vector<long long> bo, fo;
  void* slab;
  float *x, *tmp, *qkv, *ffb, *hproj;
  __nv_bfloat16* pl;           // hi / lo planes scratch
  std::vector<FePlan*> plans;
  std::vector<float*> pp_pool;   // position-projection buffers of evicted plans (each sized for max_tokens): reused, never freed
  long long launches, clock;
};

static const __nv_bfloat16* FBW(cbx_flowfe* h, int i) { return h->wbf + h->bo[i]; }
static const float* FFW(cbx_flowfe* h, int i) { return h->wf + h->fo[i]; }

// C = (A_hi + A_lo) W^T (+ bias) (+ resid), A planes packed [2][M][K] at h->pl
static int fe_gemm(cbx_flowfe* h, GemmDesc* d, int wi, int M, int N, int K, const float* bias, const float* resid, float* out) {
  GemmEpi e{};
  e.bias = bias; e.resid = resid; e.ld_resid = N; e.out_f32 = out; e.ld_f32 = N;
  // narrow column tiles + the automatic 2-way split-K: ~100 CTAs per GEMM at T ~ 400-800
  const int n_tile = (N % 64 == 0) ? 64 : 0;
  return gemm_make_hilo_ex(d, h->pl, K, (long long)M * K, FBW(h, wi), M, N, N, K, e, n_tile, 0);
}

static int fe_plan(cbx_flowfe* h, int T0, FePlan** out, cudaStream_t st) {
  for (FePlan* p : h->plans) if (p->T0 == T0) { p->stamp = ++h->clock; *out = p; return CBX_OK; }
  if ((int)h->plans.size() >= FE_MAX_PLANS) {                 // evict the least recently used length
    size_t v = 0;
    for (size_t i = 1; i < h->plans.size(); ++i) if (h->plans[i]->stamp < h->plans[v]->stamp) v = i;
    // its buffer goes back to the pool: the next plan's kernels are enqueued behind the evicted plan's last use on the
    // handle's stream, so no synchronisation (and no cudaFree / cudaMalloc: a new length used to cost 6 ms of driver time)
    h->pp_pool.push_back(h->plans[v]->pp);
    delete h->plans[v];
    h->plans.erase(h->plans.begin() + v);
  }
  FePlan* p = new FePlan();
  p->T0 = T0; p->stamp = ++h->clock; p->pp = nullptr;
  const int T1 = 2 * T0, R0 = 2 * T0 - 1, R1 = 2 * T1 - 1;
  // every buffer is sized for max_tokens, so any plan can take any pooled buffer
  const size_t pp_floats = ((size_t)N_LAYERS0 * (2 * h->max_tokens - 1) + (size_t)N_LAYERS1 * (4 * h->max_tokens - 1)) * D;
  if (!h->pp_pool.empty()) { p->pp = h->pp_pool.back(); h->pp_pool.pop_back(); }
  else if (cudaMalloc(&p->pp, pp_floats * sizeof(float)) != cudaSuccess) {
    cbx_set_error("cbx_flowfe: cudaMalloc of %zu bytes for the position projections failed", pp_floats * sizeof(float));
    delete p;
    return CBX_ERR;
  }
  int rc = 0;
  rc |= fe_gemm(h, &p->embed, BF_EMBED, T0, D, D, FFW(h, F_EMBED_B), nullptr, h->tmp);
  rc |= fe_gemm(h, &p->pl1, BF_PL1, T0, D, 4 * D, FFW(h, F_PL1_B), nullptr, h->tmp);
  rc |= fe_gemm(h, &p->pl2, BF_PL2, T0, D, 3 * D, FFW(h, F_PL2_B), h->x, h->x);
  rc |= fe_gemm(h, &p->upconv, BF_UPCONV, T1, D, 5 * D, FFW(h, F_UPCONV_B), nullptr, h->tmp);
  rc |= fe_gemm(h, &p->upembed, BF_UPEMBED, T1, D, D, FFW(h, F_UPEMBED_B), nullptr, h->x);
  rc |= fe_gemm(h, &p->proj, BF_PROJ, T1, NMEL, D, FFW(h, F_PROJ_B), nullptr, h->hproj);
  for (int l = 0; l < N_LAYERS && !rc; ++l) {
    const int T = l < N_LAYERS0 ? T0 : T1, f = F_LAYER + 10 * l, w = BF_LAYER + 5 * l;
    rc |= fe_gemm(h, &p->qkv[l], w + 0, T, 3 * D, D, FFW(h, f + 2), nullptr, h->qkv);
    rc |= fe_gemm(h, &p->out[l], w + 2, T, D, D, FFW(h, f + 3), h->x, h->x);
    rc |= fe_gemm(h, &p->w1[l], w + 3, T, FF, D, FFW(h, f + 8), nullptr, h->ffb);
    rc |= fe_gemm(h, &p->w2[l], w + 4, T, D, FF, FFW(h, f + 9), h->x, h->x);
  }
  // linear_pos(pos_emb) per layer: input-independent, computed once per length (attention.py:289-291)
  size_t off = 0;
  for (int l = 0; l < N_LAYERS && !rc; ++l) {
    const int T = l < N_LAYERS0 ? T0 : T1, R = 2 * T - 1;
    if (l == 0 || l == N_LAYERS0) k_sf_posemb<<<R, 256, 0, st>>>(T, h->pl, (long long)R * D);
    GemmDesc d;
    GemmEpi e{};
    e.out_f32 = p->pp + off; e.ld_f32 = D;
    rc |= gemm_make_hilo_ex(&d, h->pl, D, (long long)R * D, FBW(h, BF_LAYER + 5 * l + 1), R, D, D, D, e, 64, 0);
    if (!rc) rc |= gemm_launch(d, st);
    off += (size_t)R * D;
  }
  if (rc || cudaGetLastError() != cudaSuccess) { h->pp_pool.push_back(p->pp); delete p; if (!rc) cbx_set_error("cbx_flowfe: plan kernels failed"); return CBX_ERR; }
  h->plans.push_back(p);
  *out = p;
  return CBX_OK;
}

extern "C" int cbx_flowfe_create(int device, int max_tokens, const CbxFlowFrontWeights* w, cbx_flowfe** out) {
  CBX_CHECK(w && out && w->bf16_blob && w->f32_blob && w->bf16_off && w->f32_off, "cbx_flowfe_create: null argument");
  CBX_CHECK(max_tokens >= 2 && max_tokens <= 8192, "cbx_flowfe_create: max_tokens %d out of range", max_tokens);
  CBX_ON_DEVICE(device);
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, device));
  CBX_CHECK(prop.major == 10, "cbx_flowfe_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  cbx_flowfe* h = new cbx_flowfe();
  h->device = device; h->max_tokens = max_tokens; h->launches = 0; h->clock = 0;
  h->wbf = static_cast<const __nv_bfloat16*>(w->bf16_blob); h->wf = w->f32_blob;
  h->bo.assign(w->bf16_off, w->bf16_off + BF_COUNT);
  h->fo.assign(w->f32_off, w->f32_off + F_COUNT);
  const size_t T1 = 2 * (size_t)max_tokens, RP = 2 * T1;                 // rows of the widest activation / of the pos-emb planes
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 1023) / 1024 * 1024; return o; };
  const size_t o_x = take(T1 * D * 4), o_tmp = take(T1 * D * 4), o_qkv = take(T1 * 3 * D * 4), o_ff = take(T1 * FF * 4);
  const size_t o_hp = take(T1 * NMEL * 4), o_pl = take(2 * (T1 * 5 * D > RP * D ? T1 * 5 * D : RP * D) * 2);
  if (cudaMalloc(&h->slab, off) != cudaSuccess) {
    cbx_set_error("cbx_flowfe_create: cudaMalloc of %zu bytes failed", off);
    delete h;
    return CBX_ERR;
  }
  uint8_t* s = static_cast<uint8_t*>(h->slab);
  h->x = (float*)(s + o_x); h->tmp = (float*)(s + o_tmp); h->qkv = (float*)(s + o_qkv); h->ffb = (float*)(s + o_ff);
  h->hproj = (float*)(s + o_hp); h->pl = (__nv_bfloat16*)(s + o_pl);
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_flowfe_destroy(cbx_flowfe* h) {
  if (!h) return CBX_OK;
  for (FePlan* p : h->plans) { cudaFree(p->pp); delete p; }
  for (float* b : h->pp_pool) cudaFree(b);
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

// one conformer block (encoder_layer.py:155-236 without macaron / conv module; pre-norm, eps 1e-12) on x [T][512]
static int fe_layer(cbx_flowfe* h, FePlan* p, int l, int T, const float* pp, cudaStream_t st) {
  const int f = F_LAYER + 10 * l;
  const long long plane = (long long)T * D;
  k_sf_ln<<<T, 128, 0, st>>>(h->x, FFW(h, f + 0), FFW(h, f + 1), 1e-12f, 1.0f, nullptr, h->pl, plane);
  if (gemm_launch(p->qkv[l], st)) return CBX_ERR;
  static const bool relattn_tc = [] { const char* v = getenv("CBX_RELATTN_TC"); return !(v && v[0] == '0'); }();
  if (relattn_tc) {
    static CbxPerDeviceOnce once;
    if (once.first()) CBX_CUDA(cudaFuncSetAttribute(k_sf_relattn_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, RT_SMEM));
    k_sf_relattn_tc<<<dim3((T + RT_Q - 1) / RT_Q, NH), 128, RT_SMEM, st>>>(h->qkv, pp, FFW(h, f + 4), FFW(h, f + 5), T, h->pl, plane);
  } else {
    k_sf_relattn<<<dim3((T + RA_Q - 1) / RA_Q, NH), 256, 0, st>>>(h->qkv, pp, FFW(h, f + 4), FFW(h, f + 5), T, h->pl, plane);
  }
  if (gemm_launch(p->out[l], st)) return CBX_ERR;                                     // x += linear_out(attn)
  k_sf_ln<<<T, 128, 0, st>>>(h->x, FFW(h, f + 6), FFW(h, f + 7), 1e-12f, 1.0f, nullptr, h->pl, plane);
  if (gemm_launch(p->w1[l], st)) return CBX_ERR;
  const long long n4 = (long long)T * FF / 4;
  k_sf_act_hilo<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(h->ffb, h->pl, (long long)T * FF, n4, 1);
  if (gemm_launch(p->w2[l], st)) return CBX_ERR;                                      // x += w_2(silu(w_1(LN(x))))
  h->launches += 8;
  return CBX_OK;
}

extern "C" int cbx_flowfe_run(cbx_flowfe* h, const int64_t* prompt_token, int Tp, const int64_t* token, int T,
                              const float* prompt_feat, int Lp, const float* embedding, float* mu_out, float* spks_out,
                              float* cond_out, void* stream) {
  CBX_CHECK(h && token && embedding && mu_out && spks_out && cond_out, "cbx_flowfe_run: null argument");
  CBX_CHECK(Tp >= 0 && T >= 1 && (Tp == 0 || prompt_token), "cbx_flowfe_run: token counts %d + %d invalid", Tp, T);
  const int T0 = Tp + T, T1 = 2 * T0;
  CBX_CHECK(T0 <= h->max_tokens, "cbx_flowfe_run: %d tokens exceed max_tokens %d", T0, h->max_tokens);
  CBX_CHECK(Lp >= 0 && Lp <= T1 && (Lp == 0 || prompt_feat), "cbx_flowfe_run: %d prompt mel frames for %d output frames", Lp, T1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->device);
  FePlan* p;
  if (fe_plan(h, T0, &p, st)) return CBX_ERR;
  const float sqrt_d = sqrtf((float)D);
  // token embedding -> embed (Linear + LayerNorm + x sqrt(512), subsampling.py:69-113, embedding.py:256-268)
  k_sf_embed<<<T0, 128, 0, st>>>(reinterpret_cast<const long long*>(prompt_token), Tp, reinterpret_cast<const long long*>(token), T,
                                 FFW(h, F_TABLE), h->pl, (long long)T0 * D);
  if (gemm_launch(p->embed, st)) return CBX_ERR;
  k_sf_ln<<<T0, 128, 0, st>>>(h->tmp, FFW(h, F_EMBED_G), FFW(h, F_EMBED_LB), 1e-5f, sqrt_d, h->x, nullptr, 0);
  // PreLookaheadLayer (upsample_encoder.py:66-96): conv k4 with 3 frames of right context, leaky_relu, causal conv k3, residual
  k_sf_im2col<<<dim3(T0, 4), 128, 0, st>>>(h->x, T0, 1, 4, 0, 0, h->pl, (long long)T0 * 4 * D);
  if (gemm_launch(p->pl1, st)) return CBX_ERR;
  k_sf_im2col<<<dim3(T0, 3), 128, 0, st>>>(h->tmp, T0, 1, 3, 2, 2, h->pl, (long long)T0 * 3 * D);
  if (gemm_launch(p->pl2, st)) return CBX_ERR;
  h->launches += 7;
  size_t off = 0;
  for (int l = 0; l < N_LAYERS0; ++l) {
    if (fe_layer(h, p, l, T0, p->pp + off, st)) return CBX_ERR;
    off += (size_t)(2 * T0 - 1) * D;
  }
  // Upsample1D (upsample_encoder.py:37-63): nearest x2 + conv k5 with 4 frames of left padding; then the second embed
  k_sf_im2col<<<dim3(T1, 5), 128, 0, st>>>(h->x, T0, 2, 5, 4, 0, h->pl, (long long)T1 * 5 * D);
  if (gemm_launch(p->upconv, st)) return CBX_ERR;
  const long long n4 = (long long)T1 * D / 4;
  k_sf_act_hilo<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(h->tmp, h->pl, (long long)T1 * D, n4, 0);
  if (gemm_launch(p->upembed, st)) return CBX_ERR;              // -> x (pre-norm)
  k_sf_ln<<<T1, 128, 0, st>>>(h->x, FFW(h, F_UPEMBED_G), FFW(h, F_UPEMBED_LB), 1e-5f, sqrt_d, h->x, nullptr, 0);
  h->launches += 5;
  for (int l = N_LAYERS0; l < N_LAYERS; ++l) {
    if (fe_layer(h, p, l, T1, p->pp + off, st)) return CBX_ERR;
    off += (size_t)(2 * T1 - 1) * D;
  }
  // after_norm, encoder_proj (flow.py:262), mu / cond in the CFM's channel-first layout, speaker projection
  k_sf_ln<<<T1, 128, 0, st>>>(h->x, FFW(h, F_AFTER_G), FFW(h, F_AFTER_B), 1e-5f, 1.0f, nullptr, h->pl, (long long)T1 * D);
  if (gemm_launch(p->proj, st)) return CBX_ERR;
  k_sf_finish<<<T1, NMEL, 0, st>>>(h->hproj, prompt_feat, Lp, T1, mu_out, cond_out);
  k_sf_spk<<<1, 96, 0, st>>>(embedding, FFW(h, F_SPK_W), FFW(h, F_SPK_B), spks_out);
  CBX_CUDA(cudaGetLastError());
  h->launches += 4;
  return CBX_OK;
}

extern "C" int cbx_flowfe_launches(cbx_flowfe* h, int64_t* launches) {
  CBX_CHECK(h && launches, "cbx_flowfe_launches: null argument");
  *launches = h->launches;
  return CBX_OK;
}
