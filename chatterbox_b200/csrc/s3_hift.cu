// HiFT-GAN vocoder on sm_100a (reference src/chatterbox/models/s3gen/hifigan.py:291-474, f0_predictor.py:19-55):
// f0 predictor -> harmonic sine source -> STFT(16,4) -> conv_pre -> 3 x {ConvTranspose1d, source fusion,
// 3 Snake ResBlocks averaged} -> conv_post -> exp/sin -> iSTFT(16,4) -> clamp (-> trim_fade, s3gen.py:296).
// Every convolution with >= 64 input channels runs on the tcgen05 GEMM of s3_gemm.cu: dilated taps are
// row-shifted TMA loads of the channel-last activation, transposed convolutions are polyphase GEMMs whose
// N dimension enumerates (phase, channel) so the output lands directly in [time][channel] order, Snake /
// ELU / leaky-ReLU / residual adds live in the epilogues.  The 18-channel source convolutions, the 16-point
// STFT / iSTFT and the sine generator are small element-wise kernels.
#include "s3_gemm.cuh"
#include "../../include/cbx.h"
#include <map>
#include <vector>

using namespace cbx;

namespace {

constexpr int NMEL = 80, MELP = 128, F0C = 512, SR = 24000, UPS = 480, NHARM = 9;
constexpr int ST_C[3] = {256, 128, 64};
constexpr int ST_UP[3] = {8, 5, 3}, ST_UPK[3] = {16, 11, 7}, ST_UPTAPS[3] = {2, 3, 3};
constexpr int SD_K[3] = {30, 6, 1}, SD_S[3] = {15, 3, 1}, SD_P[3] = {7, 1, 0};
constexpr int RB_K[12] = {7, 7, 11, 3, 7, 11, 3, 7, 11, 3, 7, 11};   // source_resblocks 0..2, then resblocks 0..8
// blob indices (chatterbox_b200/hift.py packs in this order)
enum { B_F0 = 0, B_PRE = 5, B_UPS = 6, B_RB = 9, B_POST = 9 + 72, B_COUNT = 82 };
enum { F_F0B = 0, F_CLSW = 5, F_CLSB = 6, F_PREB = 7, F_UPB = 8, F_SDW = 11, F_SDB = 14, F_RB = 17, F_POSTB = 17 + 144,
       F_SRCW = 162, F_SRCB = 163, F_COUNT = 164 };

__device__ __forceinline__ float snake_d(float x, float a) { const float s = sinf(x * a); return x + (1.0f / (a + 1e-9f)) * s * s; }

// mel [80][L] fp32 (channel-first) -> bf16 [L][128] (channels 80..127 zero)
__global__ void k_mel_pack(const float* __restrict__ mel, __nv_bfloat16* __restrict__ out, int L) {
  const int t = blockIdx.x, c = threadIdx.x;
  out[(size_t)t * MELP + c] = __float2bfloat16_rn(c < NMEL ? mel[(size_t)c * L + t] : 0.f);
}
// f0[t] = | w . x[t] + b |   (f0_predictor.py:52-55)
__global__ void k_f0_head(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                          float* __restrict__ f0, int L) {
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (t >= L) return;
  float acc = 0.f;
  for (int k = lane; k < F0C; k += 32) acc += __bfloat162float(x[(size_t)t * F0C + k]) * w[k];
  acc = warp_sum(acc);
  if (lane == 0) f0[t] = fabsf(acc + b[0]);
}
// exclusive prefix over frames of 480 * f0 (double): phase continuity of the sine source
__global__ void k_f0_prefix(const float* __restrict__ f0, double* __restrict__ prefix, int L) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double acc = 0.0;
    for (int t = 0; t < L; ++t) { prefix[t] = acc; acc += (double)UPS * (double)f0[t]; }
  }
}
__device__ __forceinline__ void philox2(unsigned long long seed, unsigned long long idx, float& n0, float& n1) {
  unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
  unsigned c0 = (unsigned)idx, c1 = (unsigned)(idx >> 32), c2 = 0x2545F491u, c3 = 0x9E3779B9u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const unsigned a0 = hi1 ^ c1 ^ k0, a1 = lo1, a2 = hi0 ^ c3 ^ k1, a3 = lo0;
    c0 = a0; c1 = a1; c2 = a2; c3 = a3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  const float u0 = ((c0 >> 8) + 0.5f) * (1.0f / 16777216.0f), u1 = (c1 >> 8) * (1.0f / 16777216.0f);
  const float r = sqrtf(-2.0f * logf(u0));
  n0 = r * cosf(6.283185307179586f * u1);
  n1 = r * sinf(6.283185307179586f * u1);
}
// SineGen + SourceModuleHnNSF (hifigan.py:200-231, 267-283): s[n] = tanh(b + sum_h w_h (0.1 sin(theta_h + phi_h) uv + amp * noise_h))
// phase [9] and noise [9][T] may be injected (parity tests); otherwise drawn from Philox(seed).
__global__ void k_source(const float* __restrict__ f0, const double* __restrict__ prefix, const float* __restrict__ phase_in,
                         const float* __restrict__ noise_in, unsigned long long seed, const float* __restrict__ lw,
                         const float* __restrict__ lb, float* __restrict__ s, int T) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= T) return;
  const int f = n / UPS, i = n % UPS;
  const float f0v = f0[f];
  const float uv = f0v > 10.0f ? 1.f : 0.f;
  const float namp = uv * 0.003f + (1.f - uv) * (0.1f / 3.f);
  const double cum = prefix[f] + (double)(i + 1) * (double)f0v;
  float acc = lb[0];
#pragma unroll
  for (int h = 0; h < NHARM; ++h) {
    double c = cum * (double)(h + 1) / (double)SR;
    c -= floor(c);
    float ph;
    if (phase_in) ph = phase_in[h];
    else { float a, b; philox2(seed, 0x7000000000000000ull + h, a, b); ph = (h == 0) ? 0.f : 3.14159265f * erff(a * 0.70710678f); }   // U(-pi, pi)
    float nz;
    if (noise_in) nz = noise_in[(size_t)h * T + n];
    else { float a, b; philox2(seed, (unsigned long long)h * (unsigned long long)T + n, a, b); nz = a; }
    const float sw = 0.1f * sinf((float)(6.283185307179586 * c) + ph) * uv + namp * nz;
    acc += lw[h] * sw;
  }
  s[n] = tanhf(acc);
}
// torch.stft(s, 16, 4, 16, hann periodic, center=True reflect): out [T/4+1][18] = 9 real then 9 imaginary parts
__global__ void k_stft(const float* __restrict__ s, float* __restrict__ out, int T, int frames) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= frames) return;
  float x[16];
#pragma unroll
  for (int n = 0; n < 16; ++n) {
    int idx = 4 * m + n - 8;
    if (idx < 0) idx = -idx;
    if (idx >= T) idx = 2 * (T - 1) - idx;
    const float w = 0.5f - 0.5f * cospif(2.0f * n / 16.0f);
    x[n] = w * s[idx];
  }
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    float re = 0.f, im = 0.f;
#pragma unroll
    for (int n = 0; n < 16; ++n) {
      float sn, cs;
      sincospif(2.0f * ((k * n) & 15) / 16.0f, &sn, &cs);
      re += x[n] * cs;
      im -= x[n] * sn;
    }
    out[(size_t)m * 18 + k] = re;
    out[(size_t)m * 18 + 9 + k] = im;
  }
}
// source_downs[i]: Conv1d(18 -> C, k, stride, pad) on s_stft [frames][18]; w packed [18*k][C].  A CTA computes SD_TT
// consecutive output frames for all C channels (thread = channel): the input window is staged in shared memory once and every
// weight read from global memory feeds SD_TT FMAs (one frame per CTA re-read the whole 18 k x C weight matrix per frame:
// 120 us per launch at 512 mel frames).  Per output the taps are added in the same order as before (j outer, ci inner;
// frames outside the signal contribute exact zeros), so the values are unchanged.
constexpr int SD_TT = 8;
__global__ void k_source_down(const float* __restrict__ sst, const float* __restrict__ w, const float* __restrict__ b,
                              const float* __restrict__ alpha, float* __restrict__ out, __nv_bfloat16* __restrict__ out_sn,
                              int frames, int T, int Cc, int k, int stride, int pad) {
  extern __shared__ float sd_win[];            // [(SD_TT - 1) stride + k][18]
  const int t0 = blockIdx.x * SD_TT, co = threadIdx.x;
  const int nrow = (SD_TT - 1) * stride + k, src0 = t0 * stride - pad;
  for (int i = threadIdx.x; i < nrow * 18; i += blockDim.x) {
    const int src = src0 + i / 18;
    sd_win[i] = (src >= 0 && src < frames) ? sst[(size_t)src * 18 + i % 18] : 0.f;
  }
  __syncthreads();
  float acc[SD_TT];
  const float bias = b[co];
#pragma unroll
  for (int tt = 0; tt < SD_TT; ++tt) acc[tt] = bias;
  for (int j = 0; j < k; ++j) {
#pragma unroll
    for (int ci = 0; ci < 18; ++ci) {
      const float wv = w[(size_t)(ci * k + j) * Cc + co];
#pragma unroll
      for (int tt = 0; tt < SD_TT; ++tt) acc[tt] += wv * sd_win[(tt * stride + j) * 18 + ci];
    }
  }
  const float al = alpha[co];
#pragma unroll
  for (int tt = 0; tt < SD_TT; ++tt) {
    const int t = t0 + tt;
    if (t < T) {
      out[(size_t)t * Cc + co] = acc[tt];
      out_sn[(size_t)t * Cc + co] = __float2bfloat16_rn(snake_d(acc[tt], al));
    }
  }
}
// x = pad(x_up) + si  (hifigan.py:423-429; reflect == 1 is ReflectionPad1d((1,0)) of the last stage);
// also emits Snake(x) with the first activation of each of the three parallel ResBlocks
__global__ void k_fuse(const float* __restrict__ xup, const float* __restrict__ si, const float* __restrict__ a0,
                       const float* __restrict__ a1, const float* __restrict__ a2, float* __restrict__ x,
                       __nv_bfloat16* __restrict__ o0, __nv_bfloat16* __restrict__ o1, __nv_bfloat16* __restrict__ o2,
                       int T, int Cc, int reflect) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)T * Cc) return;
  const int t = (int)(idx / Cc), c = (int)(idx % Cc);
  const int tt = reflect ? (t == 0 ? 1 : t - 1) : t;
  const float v = xup[(size_t)tt * Cc + c] + si[idx];
  x[idx] = v;
  o0[idx] = __float2bfloat16_rn(snake_d(v, a0[c]));
  o1[idx] = __float2bfloat16_rn(snake_d(v, a1[c]));
  o2[idx] = __float2bfloat16_rn(snake_d(v, a2[c]));
}
// xs / 3 -> leaky_relu(slope) -> bf16   (hifigan.py:431-437, 415, 439)
__global__ void k_mean_lrelu(const float* __restrict__ acc, __nv_bfloat16* __restrict__ out, size_t n, float slope) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  float v = acc[idx] * (1.0f / 3.0f);
  v = v > 0.f ? v : v * slope;
  out[idx] = __float2bfloat16_rn(v);
}
// conv_post output [frames][32] -> magnitude/phase -> torch.istft(16, 4, hann) -> clamp -> optional trim_fade
__global__ void k_istft(const float* __restrict__ post, float* __restrict__ wav, int frames, int T, int fade) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= T) return;
  const int p = n + 8;
  int m_lo = (p - 15 + 3) / 4; if (p - 15 < 0) m_lo = 0;
  int m_hi = p / 4; if (m_hi > frames - 1) m_hi = frames - 1;
  float y = 0.f, env = 0.f;
  for (int m = m_lo; m <= m_hi; ++m) {
    const int j = p - 4 * m;
    const float w = 0.5f - 0.5f * cospif(2.0f * j / 16.0f);
    const float* x = post + (size_t)m * 32;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float mag = fminf(expf(x[k]), 100.0f);
      const float ph = sinf(x[9 + k]);
      float sp, cp;
      sincosf(ph, &sp, &cp);
      const float re = mag * cp, im = mag * sp;
      float sn, cs;
      sincospif(2.0f * ((k * j) & 15) / 16.0f, &sn, &cs);
      if (k == 0 || k == 8) acc += re * cs;
      else acc += 2.0f * (re * cs - im * sn);
    }
    y += w * acc * (1.0f / 16.0f);
    env += w * w;
  }
  float v = y / env;
  v = fminf(fmaxf(v, -0.99f), 0.99f);
  if (fade) {
    if (n < 480) v = 0.f;
    else if (n < 960) v *= 0.5f * (cosf(3.14159265358979f - (float)(n - 480) * (3.14159265358979f / 479.0f)) + 1.0f);
  }
  wav[n] = v;
}

struct HOp { int kind; GemmDesc g; int a, b; };
struct HPlan { int L; std::vector<HOp> ops; };

}  // namespace

struct cbx_hift {
  int device, max_frames;
  const __nv_bfloat16* wbf; const float* wf;
  std::vector<long long> bo, fo;
  void* slab;
  __nv_bfloat16 *mel_bf, *f0a, *f0b, *pre_bf, *xa[3], *tb, *xn, *nxt;
  float *f0, *s, *sst, *post, *xup, *si, *x, *xr, *acc;
  double* prefix;
  std::map<int, HPlan> plans;
  long long launches;
};

static const __nv_bfloat16* HB(cbx_hift* h, int i) { return h->wbf + h->bo[i]; }
static const float* HF(cbx_hift* h, int i) { return h->wf + h->fo[i]; }

enum { HK_GEMM = 0, HK_SRC_DOWN = 1, HK_FUSE = 2, HK_MEAN = 3 };

static int hift_plan(cbx_hift* h, int L, HPlan** out) {
  auto it = h->plans.find(L);
  if (it != h->plans.end()) { *out = &it->second; return CBX_OK; }
  HPlan pl; pl.L = L;
  int rc = CBX_OK;
  auto G = [&](const __nv_bfloat16* A, int lda, int rows_in, int wi, int N, int K, int taps, int dil, int pad, int M_out,
               const GemmEpi& e) {
    HOp op; op.kind = HK_GEMM; op.a = op.b = 0;
    if (gemm_make_conv(&op.g, A, lda, HB(h, wi), 1, rows_in, N, K, taps, dil, pad, M_out, e)) rc = CBX_ERR;
    pl.ops.push_back(op);
  };
  auto E = [&](int kind, int a, int b) { HOp op; op.kind = kind; op.a = a; op.b = b; pl.ops.push_back(op); };
  // ResBlock rbi on [T][Cc]: in_sn = Snake(activations1[0])(x) in bf16, x fp32; mode 0: result -> h->si (source branch),
  // mode 1: result -> h->acc (first of the three parallel blocks), mode 2: result += h->acc
  auto resblock = [&](int rbi, int Cc, int T, const __nv_bfloat16* in_sn, const float* x, int mode) {
    const int k = RB_K[rbi], fb = F_RB + rbi * 12, wb = B_RB + rbi * 6;
    for (int idx = 0; idx < 3; ++idx) {
      const int d = 1 + 2 * idx;
      GemmEpi e1{}; e1.bias = HF(h, fb + idx); e1.snake_alpha = HF(h, fb + 9 + idx); e1.out_bf16 = h->tb; e1.ld_bf16 = Cc;
      G(idx == 0 ? in_sn : h->xn, Cc, T, wb + idx, Cc, Cc, k, d, d * (k - 1) / 2, T, e1);
      GemmEpi e2{}; e2.bias = HF(h, fb + 3 + idx); e2.resid = (idx == 0) ? x : h->xr; e2.ld_resid = Cc;
      if (idx < 2) {
        e2.out_f32 = h->xr; e2.ld_f32 = Cc;
        e2.out_bf16 = h->xn; e2.ld_bf16 = Cc; e2.snake_alpha = HF(h, fb + 6 + idx + 1);
      } else if (mode == 0) {
        e2.out_f32 = h->si; e2.ld_f32 = Cc;
      } else {
        if (mode == 2) { e2.resid2 = h->acc; e2.ld_resid2 = Cc; }
        e2.out_f32 = h->acc; e2.ld_f32 = Cc;
      }
      G(h->tb, Cc, T, wb + 3 + idx, Cc, Cc, k, 1, (k - 1) / 2, T, e2);
    }
  };
  // conv_pre (80 -> 512, k7) with the leaky_relu(0.1) that precedes ups[0]
  { GemmEpi e{}; e.bias = HF(h, F_PREB); e.lrelu = 0.1f; e.out_bf16 = h->pre_bf; e.ld_bf16 = 512;
    G(h->mel_bf, MELP, L, B_PRE, 512, MELP, 7, 1, 3, L, e); }
  int T_in = L;
  const __nv_bfloat16* up_in = h->pre_bf;
  int c_in = 512;
  for (int i = 0; i < 3; ++i) {
    const int Cc = ST_C[i], u = ST_UP[i], p = (ST_UPK[i] - u) / 2;
    const int T_up = T_in * u, T = (i == 2) ? T_up + 1 : T_up;
    // source branch: source_downs[i] -> source_resblocks[i]
    E(HK_SRC_DOWN, i, T);
    resblock(i, Cc, T, h->xa[0], h->si, 0);
    // ups[i] as a polyphase GEMM: N enumerates (phase, channel); rows q = 0..T_in
    { GemmEpi e{}; e.bias = HF(h, F_UPB + i); e.bias_mod = Cc; e.out_f32 = h->xup; e.ld_f32 = u * Cc;
      e.out_shift = -(long long)p * Cc; e.out_limit = (long long)T_up * Cc;
      G(up_in, c_in, T_in, B_UPS + i, u * Cc, c_in, ST_UPTAPS[i], -1, 0, T_in + 1, e); }
    E(HK_FUSE, i, T);
    for (int j = 0; j < 3; ++j) resblock(3 + i * 3 + j, Cc, T, h->xa[j], h->x, j == 0 ? 1 : 2);
    E(HK_MEAN, i, T);
    T_in = T; up_in = h->nxt; c_in = Cc;
  }
  { GemmEpi e{}; e.bias = HF(h, F_POSTB); e.out_f32 = h->post; e.ld_f32 = 32;
    G(h->nxt, 64, T_in, B_POST, 32, 64, 7, 1, 3, T_in, e); }
  if (rc) return CBX_ERR;
  auto ins = h->plans.emplace(L, std::move(pl));
  *out = &ins.first->second;
  return CBX_OK;
}

static int hift_f0(cbx_hift* h, const float* mel, int L, cudaStream_t st) {
  k_mel_pack<<<L, MELP, 0, st>>>(mel, h->mel_bf, L);
  for (int j = 0; j < 5; ++j) {
    GemmDesc g;
    GemmEpi e{}; e.bias = HF(h, F_F0B + j); e.elu = 1; e.out_bf16 = (j & 1) ? h->f0b : h->f0a; e.ld_bf16 = F0C;
    const __nv_bfloat16* A = (j == 0) ? h->mel_bf : ((j & 1) ? h->f0a : h->f0b);
    const int K = (j == 0) ? MELP : F0C;
    if (gemm_make_conv(&g, A, K, HB(h, B_F0 + j), 1, L, F0C, K, 3, 1, 1, L, e)) return CBX_ERR;
    if (gemm_launch(g, st)) return CBX_ERR;
  }
  k_f0_head<<<(L + 7) / 8, 256, 0, st>>>(h->f0a, HF(h, F_CLSW), HF(h, F_CLSB), h->f0, L);   // j = 4 wrote f0a
  CBX_CUDA(cudaGetLastError());
  h->launches += 7;
  return CBX_OK;
}

static int hift_source(cbx_hift* h, const float* f0, int L, const float* phase, const float* noise, unsigned long long seed,
                       float* s_out, cudaStream_t st) {
  const int T = UPS * L;
  k_f0_prefix<<<1, 32, 0, st>>>(f0, h->prefix, L);
  k_source<<<(T + 255) / 256, 256, 0, st>>>(f0, h->prefix, phase, noise, seed, HF(h, F_SRCW), HF(h, F_SRCB), s_out, T);
  CBX_CUDA(cudaGetLastError());
  h->launches += 2;
  return CBX_OK;
}

static int hift_decode(cbx_hift* h, const float* mel, int L, const float* s, int fade, float* wav, cudaStream_t st) {
  HPlan* pl;
  if (hift_plan(h, L, &pl)) return CBX_ERR;
  const int T = UPS * L, frames = T / 4 + 1;
  k_mel_pack<<<L, MELP, 0, st>>>(mel, h->mel_bf, L);
  k_stft<<<(frames + 127) / 128, 128, 0, st>>>(s, h->sst, T, frames);
  h->launches += 2;
  for (HOp& op : pl->ops) {
    if (op.kind == HK_GEMM) {
      if (gemm_launch(op.g, st)) return CBX_ERR;
    } else if (op.kind == HK_SRC_DOWN) {
      const int i = op.a, Tt = op.b, Cc = ST_C[i];
      k_source_down<<<(Tt + SD_TT - 1) / SD_TT, Cc, ((SD_TT - 1) * SD_S[i] + SD_K[i]) * 18 * sizeof(float), st>>>(h->sst, HF(h, F_SDW + i), HF(h, F_SDB + i), HF(h, F_RB + i * 12 + 6), h->si, h->xa[0],
                                       frames, Tt, Cc, SD_K[i], SD_S[i], SD_P[i]);
    } else if (op.kind == HK_FUSE) {
      const int i = op.a, Tt = op.b, Cc = ST_C[i];
      const size_t n = (size_t)Tt * Cc;
      k_fuse<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->xup, h->si, HF(h, F_RB + (3 + i * 3 + 0) * 12 + 6),
                                                          HF(h, F_RB + (3 + i * 3 + 1) * 12 + 6), HF(h, F_RB + (3 + i * 3 + 2) * 12 + 6),
                                                          h->x, h->xa[0], h->xa[1], h->xa[2], Tt, Cc, i == 2);
    } else {
      const int i = op.a, Tt = op.b, Cc = ST_C[i];
      const size_t n = (size_t)Tt * Cc;
      k_mean_lrelu<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->acc, h->nxt, n, i == 2 ? 0.01f : 0.1f);
    }
    h->launches += 1;
  }
  k_istft<<<(T + 255) / 256, 256, 0, st>>>(h->post, wav, frames, T, fade);
  CBX_CUDA(cudaGetLastError());
  h->launches += 1;
  return CBX_OK;
}

// =================================================================================================
// C-ABI (include/cbx.h)
// =================================================================================================
extern "C" int cbx_hift_create(int device, int max_frames, const CbxHiftWeights* w, cbx_hift** out) {
  CBX_CHECK(w && out && w->bf16_blob && w->f32_blob && w->bf16_off && w->f32_off, "cbx_hift_create: null argument");
  CBX_CHECK(max_frames >= 1, "cbx_hift_create: max_frames must be positive");
  CBX_ON_DEVICE(device);
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, device));
  CBX_CHECK(prop.major == 10, "cbx_hift_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  cbx_hift* h = new cbx_hift();
  h->device = device; h->max_frames = max_frames; h->launches = 0;
  h->wbf = static_cast<const __nv_bfloat16*>(w->bf16_blob); h->wf = w->f32_blob;
  h->bo.assign(w->bf16_off, w->bf16_off + B_COUNT);
  h->fo.assign(w->f32_off, w->f32_off + F_COUNT);
  const size_t L = max_frames, T = (size_t)UPS * L, frames = T / 4 + 1, E = 7680 * L + 64;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 1023) / 1024 * 1024; return o; };
  const size_t o_mel = take(L * MELP * 2), o_f0a = take(L * F0C * 2), o_f0b = take(L * F0C * 2), o_pre = take(L * 512 * 2);
  const size_t o_xa0 = take(E * 2), o_xa1 = take(E * 2), o_xa2 = take(E * 2), o_tb = take(E * 2), o_xn = take(E * 2), o_nxt = take(E * 2);
  const size_t o_f0 = take(L * 4), o_s = take(T * 4), o_sst = take(frames * 18 * 4), o_post = take(frames * 32 * 4);
  const size_t o_xup = take(E * 4), o_si = take(E * 4), o_x = take(E * 4), o_xr = take(E * 4), o_acc = take(E * 4);
  const size_t o_pref = take(L * 8);
  if (cudaMalloc(&h->slab, off) != cudaSuccess) {
    cbx_set_error("cbx_hift_create: cudaMalloc of %zu bytes failed", off);
    delete h;
    return CBX_ERR;
  }
  cudaMemset(h->slab, 0, off);
  uint8_t* s = static_cast<uint8_t*>(h->slab);
  h->mel_bf = (__nv_bfloat16*)(s + o_mel); h->f0a = (__nv_bfloat16*)(s + o_f0a); h->f0b = (__nv_bfloat16*)(s + o_f0b);
  h->pre_bf = (__nv_bfloat16*)(s + o_pre);
  h->xa[0] = (__nv_bfloat16*)(s + o_xa0); h->xa[1] = (__nv_bfloat16*)(s + o_xa1); h->xa[2] = (__nv_bfloat16*)(s + o_xa2);
  h->tb = (__nv_bfloat16*)(s + o_tb); h->xn = (__nv_bfloat16*)(s + o_xn); h->nxt = (__nv_bfloat16*)(s + o_nxt);
  h->f0 = (float*)(s + o_f0); h->s = (float*)(s + o_s); h->sst = (float*)(s + o_sst); h->post = (float*)(s + o_post);
  h->xup = (float*)(s + o_xup); h->si = (float*)(s + o_si); h->x = (float*)(s + o_x); h->xr = (float*)(s + o_xr);
  h->acc = (float*)(s + o_acc); h->prefix = (double*)(s + o_pref);
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_hift_destroy(cbx_hift* h) {
  if (!h) return CBX_OK;
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

#define HIFT_ENTER(name)                                                                            \
  CBX_CHECK(h, name ": null handle");                                                               \
  CBX_CHECK(L >= 1 && L <= h->max_frames, name ": L=%d exceeds max_frames %d", L, h->max_frames);  \
  cudaStream_t st = static_cast<cudaStream_t>(stream);                                             \
  CBX_ON_DEVICE(h->device);

extern "C" int cbx_hift_f0(cbx_hift* h, const float* mel, int L, float* f0_out, void* stream) {
  HIFT_ENTER("cbx_hift_f0");
  CBX_CHECK(mel && f0_out, "cbx_hift_f0: null argument");
  if (hift_f0(h, mel, L, st)) return CBX_ERR;
  CBX_CUDA(cudaMemcpyAsync(f0_out, h->f0, (size_t)L * 4, cudaMemcpyDeviceToDevice, st));
  return CBX_OK;
}

extern "C" int cbx_hift_source(cbx_hift* h, const float* f0, int L, const float* phase, const float* noise, uint64_t seed,
                               float* s_out, void* stream) {
  HIFT_ENTER("cbx_hift_source");
  CBX_CHECK(f0 && s_out, "cbx_hift_source: null argument");
  return hift_source(h, f0, L, phase, noise, seed, s_out, st);
}

extern "C" int cbx_hift_decode(cbx_hift* h, const float* mel, int L, const float* s, int apply_fade, float* wav_out,
                               void* stream) {
  HIFT_ENTER("cbx_hift_decode");
  CBX_CHECK(mel && s && wav_out, "cbx_hift_decode: null argument");
  return hift_decode(h, mel, L, s, apply_fade, wav_out, st);
}

extern "C" int cbx_hift_inference(cbx_hift* h, const float* mel, int L, const float* source_in, uint64_t seed, int apply_fade,
                                  float* wav_out, float* source_out, void* stream) {
  HIFT_ENTER("cbx_hift_inference");
  CBX_CHECK(mel && wav_out, "cbx_hift_inference: null argument");
  const float* s = source_in;
  if (!s) {
    if (hift_f0(h, mel, L, st)) return CBX_ERR;
    if (hift_source(h, h->f0, L, nullptr, nullptr, seed, h->s, st)) return CBX_ERR;
    s = h->s;
  }
  if (source_out) CBX_CUDA(cudaMemcpyAsync(source_out, s, (size_t)UPS * L * 4, cudaMemcpyDeviceToDevice, st));
  return hift_decode(h, mel, L, s, apply_fade, wav_out, st);
}

extern "C" int cbx_hift_launches(cbx_hift* h, int64_t* launches) {
  CBX_CHECK(h && launches, "cbx_hift_launches: null argument");
  *launches = h->launches;
  return CBX_OK;
}

// =================================================================================================
// Allocation-free variant of cbx_hift_source (include/cbx.h, section 3b).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  Same two kernels as hift_source above; what differs is where the scratch
// lives: the double-precision phase prefix (one entry per mel frame) sits in a file-scope device array instead of the
// handle's cudaMalloc'ed slab, and the two SourceModuleHnNSF linear weights come from the caller, so no handle is
// needed either.  Every GPU that loads the library holds the array (32 KB) until exit.
//   * The array is shared by every call on a device and keeps the last call's contents.  Nothing has to start at
//     zero: k_f0_prefix rewrites entries [0, L) before k_source reads them, on the same stream.
//   * Calls that overlap on the GPU (different streams, or host threads on different streams) RACE on the array:
//     serialise them on one stream.
// In host code the array's name is a host-side stand-in, not a device address: its address is taken once per device
// with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
// =================================================================================================
#define CBX_HIFT_SOURCE_STATIC_MAX_FRAMES 4096      // 4096 mel frames = 81.92 s of audio at 24 kHz
static __device__ __align__(256) double g_source_prefix_scratch[CBX_HIFT_SOURCE_STATIC_MAX_FRAMES];

extern "C" int cbx_hift_source_static_max_frames(void) { return CBX_HIFT_SOURCE_STATIC_MAX_FRAMES; }

extern "C" int cbx_hift_source_static(int device, const float* lin_w, const float* lin_b, const float* f0, int L,
                                      const float* phase, const float* noise, uint64_t seed, float* s_out, void* stream) {
  CBX_CHECK(lin_w && lin_b && f0 && s_out, "cbx_hift_source_static: null argument");
  CBX_CHECK(L >= 1, "cbx_hift_source_static: L must be >= 1 (got %d)", L);
  CBX_CHECK(L <= CBX_HIFT_SOURCE_STATIC_MAX_FRAMES,
            "cbx_hift_source_static: L=%d exceeds the static scratch (%d mel frames); use cbx_hift_source with a handle sized for it",
            L, CBX_HIFT_SOURCE_STATIC_MAX_FRAMES);
  CBX_CHECK(device >= 0 && device < 64, "cbx_hift_source_static: device ordinal %d out of range", device);
  CBX_ON_DEVICE(device);
  static double* scratch_of_device[64] = {};          // device address of the array, per device (set up once each)
  if (!scratch_of_device[device]) {
    double* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_source_prefix_scratch);
    CBX_CHECK(e == cudaSuccess && ptr, "cbx_hift_source_static: cudaGetSymbolAddress failed on device %d (%s)", device,
              cudaGetErrorString(e));
    scratch_of_device[device] = ptr;
  }
  double* prefix = scratch_of_device[device];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int T = UPS * L;
  k_f0_prefix<<<1, 32, 0, st>>>(f0, prefix, L);
  k_source<<<(T + 255) / 256, 256, 0, st>>>(f0, prefix, phase, noise, seed, lin_w, lin_b, s_out, T);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}
