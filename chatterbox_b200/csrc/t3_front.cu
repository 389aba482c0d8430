// T3 conditioning front-end for sm_100a: everything `T3.inference` does before the first transformer pass, once per
// utterance — T3CondEnc.forward (src/chatterbox/models/t3/modules/cond_enc.py:64-97: speaker projection, Perceiver
// resampler over the conditioning-prompt speech tokens, emotion projection), Perceiver / AttentionBlock2
// (modules/perceiver.py:113-212: shared LayerNorm, q/k/v/out Linear(1024, 1024) with bias, 4 heads x 256, softmax
// attention, run twice: cross then self), T3.prepare_input_embeds (t3.py:80-108: text / speech embeddings + learned
// position embeddings, zeroed text embeddings in the unconditional CFG row, concat) and the doubled BOS of t3.py:284-295.
// Output: the prompt embeddings [B][P][1024] fp32 that cbx_t3_prefill / cbx_t3_batch_prefill consume.
//
// The reference runs this in fp32 and so does this file: M <= 150 rows against 1024 x 1024 fp32 matrices is ~2 GFLOP per
// utterance in total — latency-bound, not a tensor-core problem — and fp32 keeps the prefix bit-close to the reference
// (the T3 parity tests with an fp32 KV slab sit at 2e-5).  ~16 small launches on one stream, no host synchronisation:
// the emotion scalar and all token ids are read on the device.
#include "common.cuh"
#include "../../include/cbx.h"

namespace {

using namespace cbx;

constexpr int H = 1024, NQ = 32, NHEAD = 4, HDIM = 256, SPK = 256;
constexpr int START_SPEECH = 6561;
constexpr float LN_EPS = 1e-5f;

// X[r] = speech_emb[tok[r]] + speech_pos_emb[r]        (t3.py:73-76)
__global__ void __launch_bounds__(256) k_fe_gather(const long long* __restrict__ tok, const float* __restrict__ emb,
                                                   const float* __restrict__ pos, float* __restrict__ out) {
  const int r = blockIdx.x, t = threadIdx.x;
  const float4 a = __ldg(reinterpret_cast<const float4*>(emb + (size_t)tok[r] * H) + t);
  const float4 b = __ldg(reinterpret_cast<const float4*>(pos + (size_t)r * H) + t);
  reinterpret_cast<float4*>(out + (size_t)r * H)[t] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

// nn.LayerNorm(1024) over a row (biased variance, eps 1e-5): one row per CTA
__global__ void __launch_bounds__(256) k_fe_ln(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b,
                                               float* __restrict__ out) {
  __shared__ float red[2][8];
  const int r = blockIdx.x, t = threadIdx.x;
  const float4 v = reinterpret_cast<const float4*>(x + (size_t)r * H)[t];
  float s = warp_sum(v.x + v.y + v.z + v.w);
  if ((t & 31) == 0) red[0][t >> 5] = s;
  __syncthreads();
  float mean = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) mean += red[0][w];
  mean *= 1.0f / H;
  const float dx = v.x - mean, dy = v.y - mean, dz = v.z - mean, dw = v.w - mean;
  float q = warp_sum(dx * dx + dy * dy + dz * dz + dw * dw);
  if ((t & 31) == 0) red[1][t >> 5] = q;
  __syncthreads();
  float var = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) var += red[1][w];
  const float rstd = rsqrtf(var * (1.0f / H) + LN_EPS);
  const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + t), bb = __ldg(reinterpret_cast<const float4*>(b) + t);
  reinterpret_cast<float4*>(out + (size_t)r * H)[t] =
      make_float4(dx * rstd * gg.x + bb.x, dy * rstd * gg.y + bb.y, dz * rstd * gg.z + bb.z, dw * rstd * gg.w + bb.w);
}

// C[M][N] = A[M][K] W[N][K]^T + bias[N] (+ resid[M][N]), fp32 on CUDA cores.  CTA = 32 rows x 32 columns, 256 threads,
// thread = 2 rows x 2 columns.  K is walked in tiles of 128: per tile every thread has 8 independent 128-bit global loads in
// flight (the loop is bound by load latency, not arithmetic: M <= 160), then 128 steps from shared memory (odd row stride:
// the 32 weight rows a warp reads at one k fall into 32 different banks; its 2 activation rows are broadcasts).
constexpr int SG_BM = 32, SG_BN = 32, SG_BK = 128, SG_LD = SG_BK + 1;
__global__ void __launch_bounds__(256) k_fe_sgemm(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                                                  const float* __restrict__ resid, float* __restrict__ C, int M, int N, int K) {
  __shared__ float As[SG_BM * SG_LD], Ws[SG_BN * SG_LD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  float acc[2][2] = {};
  for (int k0 = 0; k0 < K; k0 += SG_BK) {
    float4 av[4], wv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {              // 32 rows x 32 float4 per operand = 1024 float4 = 4 per thread
      const int idx = tid + i * 256, r = idx >> 5, c4 = idx & 31;
      av[i] = (m0 + r < M) ? *reinterpret_cast<const float4*>(A + (size_t)(m0 + r) * K + k0 + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      wv[i] = __ldg(reinterpret_cast<const float4*>(W + (size_t)(n0 + r) * K + k0 + c4 * 4));
    }
    __syncthreads();                           // the previous tile's readers are done
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256, r = idx >> 5, c4 = idx & 31;
      float* da = As + r * SG_LD + c4 * 4;
      float* dw = Ws + r * SG_LD + c4 * 4;
      da[0] = av[i].x; da[1] = av[i].y; da[2] = av[i].z; da[3] = av[i].w;
      dw[0] = wv[i].x; dw[1] = wv[i].y; dw[2] = wv[i].z; dw[3] = wv[i].w;
    }
    __syncthreads();
    const float* a0 = As + (ty * 2) * SG_LD;
    const float* w0 = Ws + (tx * 2) * SG_LD;
#pragma unroll 16
    for (int k = 0; k < SG_BK; ++k) {
      const float x0 = a0[k], x1 = a0[SG_LD + k], y0 = w0[k], y1 = w0[SG_LD + k];
      acc[0][0] += x0 * y0; acc[0][1] += x0 * y1;
      acc[1][0] += x1 * y0; acc[1][1] += x1 * y1;
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int m = m0 + ty * 2 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int n = n0 + tx * 2 + j;
      float v = acc[i][j] + (bias ? __ldg(bias + n) : 0.f);
      if (resid) v += resid[(size_t)m * N + n];
      C[(size_t)m * N + n] = v;
    }
  }
}

// softmax(q k^T / 16) v for one (query, head): AttentionQKV (perceiver.py:84-100), 4 heads x 256.  256 threads: a thread
// scores key `tid` (nk <= 256), then owns output dimension `tid`.
__global__ void __launch_bounds__(256) k_fe_attn(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                                                 float* __restrict__ out, int nk) {
  __shared__ float qs[HDIM], ps[256], red[8];
  const int i = blockIdx.x, h = blockIdx.y, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  qs[t] = q[(size_t)i * H + h * HDIM + t];
  __syncthreads();
  float s = -INFINITY;
  if (t < nk) {
    const float4* kr = reinterpret_cast<const float4*>(k + (size_t)t * H + h * HDIM);
    float a = 0.f;
#pragma unroll 8
    for (int d = 0; d < HDIM / 4; ++d) {
      const float4 kk = kr[d];
      a += qs[4 * d] * kk.x + qs[4 * d + 1] * kk.y + qs[4 * d + 2] * kk.z + qs[4 * d + 3] * kk.w;
    }
    s = a * 0.0625f;          // head_dim ** -0.5
  }
  float m = warp_max(s);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
  const float e = (t < nk) ? expf(s - m) : 0.f;
  ps[t] = e;
  float l = warp_sum(e);
  __syncthreads();
  if (lane == 0) red[warp] = l;
  __syncthreads();
  l = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) l += red[w];
  float o = 0.f;
  for (int j = 0; j < nk; ++j) o += ps[j] * v[(size_t)j * H + h * HDIM + t];
  out[(size_t)i * H + h * HDIM + t] = o / l;
}

struct FeAssemble {
  const float *spk_w, *spk_b, *speaker, *per, *emo_w, *emotion, *text_emb, *text_pos, *speech_emb, *speech_pos;
  const long long* text;       // [B][N]
  int B, N, n_per, n_bos, zero_uncond_text;
  float* out;                  // [B][P][1024]
};
// one CTA per (position, batch row): speaker | perceiver rows | emotion | text | BOS (| BOS)      (cond_enc.py:90-97, t3.py:80-108)
__global__ void __launch_bounds__(256) k_fe_assemble(const FeAssemble a) {
  __shared__ float es[SPK];
  const int p = blockIdx.x, b = blockIdx.y, t = threadIdx.x;
  const int P = 1 + a.n_per + 1 + a.N + a.n_bos;
  float4* dst = reinterpret_cast<float4*>(a.out + ((size_t)b * P + p) * H) + t;
  if (p == 0) {                                              // spkr_enc: Linear(256 -> 1024) + bias
    es[t] = a.speaker[t];
    __syncthreads();
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4* wr = reinterpret_cast<const float4*>(a.spk_w + (size_t)(4 * t + j) * SPK);
      float s = 0.f;
      for (int kk = 0; kk < SPK / 4; ++kk) {
        const float4 w = __ldg(wr + kk);
        s += w.x * es[4 * kk] + w.y * es[4 * kk + 1] + w.z * es[4 * kk + 2] + w.w * es[4 * kk + 3];
      }
      o[j] = s + __ldg(a.spk_b + 4 * t + j);
    }
    *dst = make_float4(o[0], o[1], o[2], o[3]);
  } else if (p <= a.n_per) {
    *dst = reinterpret_cast<const float4*>(a.per + (size_t)(p - 1) * H)[t];
  } else if (p == a.n_per + 1) {                             // emotion_adv_fc: Linear(1 -> 1024), no bias
    const float e = a.emotion[0];
    const float4 w = __ldg(reinterpret_cast<const float4*>(a.emo_w) + t);
    *dst = make_float4(w.x * e, w.y * e, w.z * e, w.w * e);
  } else if (p < a.n_per + 2 + a.N) {
    const int tt = p - (a.n_per + 2);
    const float4 ps = __ldg(reinterpret_cast<const float4*>(a.text_pos + (size_t)tt * H) + t);
    float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!(a.zero_uncond_text && b == 1)) e = __ldg(reinterpret_cast<const float4*>(a.text_emb + (size_t)a.text[(size_t)b * a.N + tt] * H) + t);
    *dst = make_float4(e.x + ps.x, e.y + ps.y, e.z + ps.z, e.w + ps.w);
  } else {                                                   // BOS at learned speech position 0
    const float4 e = __ldg(reinterpret_cast<const float4*>(a.speech_emb + (size_t)START_SPEECH * H) + t);
    const float4 ps = __ldg(reinterpret_cast<const float4*>(a.speech_pos) + t);
    *dst = make_float4(e.x + ps.x, e.y + ps.y, e.z + ps.z, e.w + ps.w);
  }
}

}  // namespace

// =================================================================================================
// C-ABI (include/cbx.h, section 1c)
// =================================================================================================
struct cbx_t3fe {
  int device, max_cond;
  CbxT3FrontWeights w;
  void* slab;
  float *x2, *x1n, *x2n, *q, *k, *v, *o, *pre, *per;
  long long launches;
};

static int fe_sgemm(const float* A, const float* W, const float* bias, const float* resid, float* C, int M, int N, int K,
                    cudaStream_t st) {
  k_fe_sgemm<<<dim3(N / SG_BN, (M + SG_BM - 1) / SG_BM), 256, 0, st>>>(A, W, bias, resid, C, M, N, K);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

extern "C" int cbx_t3_front_create(int device, int max_cond_tokens, const CbxT3FrontWeights* w, cbx_t3fe** out) {
  CBX_CHECK(w && out, "cbx_t3_front_create: null argument");
  CBX_CHECK(w->spkr_w && w->spkr_b && w->emotion_w && w->text_emb && w->text_pos_emb && w->speech_emb && w->speech_pos_emb,
            "cbx_t3_front_create: missing embedding / projection weights");
  CBX_CHECK(max_cond_tokens >= 0 && max_cond_tokens <= 256, "cbx_t3_front_create: max_cond_tokens %d out of range (0..256)", max_cond_tokens);
  CBX_CHECK(max_cond_tokens == 0 || (w->query && w->norm_g && w->norm_b && w->wq && w->bq && w->wk && w->bk && w->wv && w->bv && w->wo && w->bo),
            "cbx_t3_front_create: missing perceiver weights");
  CBX_ON_DEVICE(device);
  cbx_t3fe* h = new cbx_t3fe();
  h->device = device; h->max_cond = max_cond_tokens; h->w = *w; h->launches = 0;
  const size_t R = (size_t)(max_cond_tokens > NQ ? max_cond_tokens : NQ) * H * sizeof(float);
  if (cudaMalloc(&h->slab, 9 * R) != cudaSuccess) {
    cbx_set_error("cbx_t3_front_create: cudaMalloc of %zu bytes failed", 9 * R);
    delete h;
    return CBX_ERR;
  }
  float* s = static_cast<float*>(h->slab);
  const size_t n = R / sizeof(float);
  h->x2 = s; h->x1n = s + n; h->x2n = s + 2 * n; h->q = s + 3 * n; h->k = s + 4 * n; h->v = s + 5 * n; h->o = s + 6 * n;
  h->pre = s + 7 * n; h->per = s + 8 * n;
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_t3_front_destroy(cbx_t3fe* h) {
  if (!h) return CBX_OK;
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

extern "C" int cbx_t3_front(cbx_t3fe* h, const float* speaker_emb, const int64_t* cond_tokens, const float* cond_prompt_emb,
                            int n_cond, const float* emotion_adv, const int64_t* text_tokens, int B, int N, int zero_uncond_text,
                            int n_bos, float* embeds_out, void* stream) {
  CBX_CHECK(h && speaker_emb && emotion_adv && text_tokens && embeds_out, "cbx_t3_front: null argument");
  CBX_CHECK(B == 1 || B == 2, "cbx_t3_front: B must be 1 or 2 (got %d)", B);
  CBX_CHECK(N >= 1, "cbx_t3_front: no text tokens");
  CBX_CHECK(n_bos == 1 || n_bos == 2, "cbx_t3_front: n_bos must be 1 or 2");
  CBX_CHECK(n_cond >= 0 && n_cond <= h->max_cond, "cbx_t3_front: %d conditioning tokens exceed max_cond_tokens %d", n_cond, h->max_cond);
  CBX_CHECK(n_cond == 0 || cond_tokens || cond_prompt_emb, "cbx_t3_front: conditioning tokens (or their embeddings) missing");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->device);
  const CbxT3FrontWeights& w = h->w;
  int n_per = 0;
  if (n_cond > 0) {
    // cond_prompt_speech_emb = speech_emb(tokens) + speech_pos_emb (t3.py:73-76) unless the caller supplies it
    const float* x2 = cond_prompt_emb;
    if (!x2) {
      k_fe_gather<<<n_cond, 256, 0, st>>>(reinterpret_cast<const long long*>(cond_tokens), w.speech_emb, w.speech_pos_emb, h->x2);
      x2 = h->x2;
    }
    // Perceiver.forward (perceiver.py:199-212): pre = attn(query, h); out = attn(pre, pre)   — the SAME block twice
    k_fe_ln<<<NQ, 256, 0, st>>>(w.query, w.norm_g, w.norm_b, h->x1n);
    k_fe_ln<<<n_cond, 256, 0, st>>>(x2, w.norm_g, w.norm_b, h->x2n);
    if (fe_sgemm(h->x1n, w.wq, w.bq, nullptr, h->q, NQ, H, H, st) || fe_sgemm(h->x2n, w.wk, w.bk, nullptr, h->k, n_cond, H, H, st) ||
        fe_sgemm(h->x2n, w.wv, w.bv, nullptr, h->v, n_cond, H, H, st)) return CBX_ERR;
    k_fe_attn<<<dim3(NQ, NHEAD), 256, 0, st>>>(h->q, h->k, h->v, h->o, n_cond);
    if (fe_sgemm(h->o, w.wo, w.bo, w.query, h->pre, NQ, H, H, st)) return CBX_ERR;
    k_fe_ln<<<NQ, 256, 0, st>>>(h->pre, w.norm_g, w.norm_b, h->x1n);
    if (fe_sgemm(h->x1n, w.wq, w.bq, nullptr, h->q, NQ, H, H, st) || fe_sgemm(h->x1n, w.wk, w.bk, nullptr, h->k, NQ, H, H, st) ||
        fe_sgemm(h->x1n, w.wv, w.bv, nullptr, h->v, NQ, H, H, st)) return CBX_ERR;
    k_fe_attn<<<dim3(NQ, NHEAD), 256, 0, st>>>(h->q, h->k, h->v, h->o, NQ);
    if (fe_sgemm(h->o, w.wo, w.bo, h->pre, h->per, NQ, H, H, st)) return CBX_ERR;
    n_per = NQ;
    h->launches += 14 + (cond_prompt_emb ? 0 : 1);
  }
  FeAssemble a;
  a.spk_w = w.spkr_w; a.spk_b = w.spkr_b; a.speaker = speaker_emb; a.per = h->per; a.emo_w = w.emotion_w; a.emotion = emotion_adv;
  a.text_emb = w.text_emb; a.text_pos = w.text_pos_emb; a.speech_emb = w.speech_emb; a.speech_pos = w.speech_pos_emb;
  a.text = reinterpret_cast<const long long*>(text_tokens); a.B = B; a.N = N; a.n_per = n_per; a.n_bos = n_bos;
  a.zero_uncond_text = zero_uncond_text; a.out = embeds_out;
  const int P = 1 + n_per + 1 + N + n_bos;
  k_fe_assemble<<<dim3(P, B), 256, 0, st>>>(a);
  CBX_CUDA(cudaGetLastError());
  h->launches += 1;
  return CBX_OK;
}

extern "C" int cbx_t3_front_launches(cbx_t3fe* h, int64_t* launches) {
  CBX_CHECK(h && launches, "cbx_t3_front_launches: null argument");
  *launches = h->launches;
  return CBX_OK;
}
