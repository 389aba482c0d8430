// S3Gen conditional flow-matching solve on sm_100a: the 10-step Euler/CFG loop of
// ConditionalCFM.solve_euler (reference src/chatterbox/models/s3gen/flow_matching.py:72-123) around the
// ConditionalDecoder estimator (decoder.py:233-317), as a fixed sequence of tcgen05 GEMMs
// (s3_gemm.cu), flash attention (s3_attn.cu) and a few element-wise kernels, all on one stream.
// Activations are channel-last [2 (cond, uncond)][L][C]; the residual stream is fp32, GEMM operands bf16.
#include "s3_gemm.cuh"
#include "../../include/cbx.h"
#include <algorithm>
#include <cstdlib>
#include <map>
#include <vector>

namespace cbx {
int s3_attn_launch(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st);
int s3_attn_launch_mma(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st);
}
using namespace cbx;

namespace {

constexpr int NRES = 14, NTFM = 56, C = 256, CIN = 320, NMEL = 80;
enum { BF_RES = 0, BF_TFM = NRES * 3, BF_TAIL = BF_TFM + NTFM * 4, BF_COUNT = BF_TAIL + 4 };
enum { F_RES = 0, F_TFM = NRES * 7, F_TAIL = F_TFM + NTFM * 7, F_TIME = F_TAIL + 6, F_COUNT = F_TIME + 6 };

__device__ __forceinline__ float mish_d(float x) {
  const float sp = (x > 20.f) ? x : log1pf(expf(x));
  return x * tanhf(sp);
}

// ---- time embedding: SinusoidalPosEmb(320, scale 1000) -> Linear -> SiLU -> Linear; Mish kept for the resnet MLPs
__global__ void k_temb(const float* __restrict__ tvals, const float* __restrict__ w1, const float* __restrict__ b1,
                       const float* __restrict__ w2, const float* __restrict__ b2, float* __restrict__ mish_temb) {
  __shared__ float emb[320];
  __shared__ float h1[1024];
  const int s = blockIdx.x, n = threadIdx.x;
  const float t = tvals[s];
  if (n < 160) {
    const float f = expf((float)n * -(logf(10000.f) / 159.f));
    const float a = 1000.f * t * f;
    emb[n] = sinf(a);
    emb[160 + n] = cosf(a);
  }
  __syncthreads();
  float acc = b1[n];
  for (int k = 0; k < 320; ++k) acc += w1[(size_t)n * 320 + k] * emb[k];
  h1[n] = acc / (1.f + expf(-acc));
  __syncthreads();
  acc = b2[n];
  for (int k = 0; k < 1024; ++k) acc += w2[(size_t)n * 1024 + k] * h1[k];
  mish_temb[(size_t)s * 1024 + n] = mish_d(acc);
}
// resnet MLP: tvec[s][r][2][256] = Linear(Mish(temb))  (matcha/decoder.py:50,57), duplicated for both CFG rows
__global__ void k_tvec(const float* __restrict__ mish_temb, const float* __restrict__ w, const float* __restrict__ bias,
                       float* __restrict__ tvec) {
  const int r = blockIdx.x, s = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* x = mish_temb + (size_t)s * 1024;
  for (int n = warp; n < C; n += blockDim.x / 32) {
    const float* wr = w + ((size_t)r * C + n) * 1024;
    float acc = 0.f;
    for (int k = lane; k < 1024; k += 32) acc += wr[k] * x[k];
    acc = warp_sum(acc);
    if (lane == 0) {
      const float v = acc + bias[r * C + n];
      float* o = tvec + ((size_t)(s * NRES + r) * 2) * C;
      o[n] = v;
      o[C + n] = v;
    }
  }
}
// xin[b][t][0:80] = x, [80:160] = mu, [160:240] = spks, [240:320] = cond (decoder.py:254-260); inputs are the
// reference's channel-first fp32 tensors per CFG row; a null row-1 pointer means zeros (flow_matching.py:95-108)
struct PackIn { const float *x[2], *mu[2], *spks[2], *cond[2]; };
__global__ void k_pack(PackIn in, float* __restrict__ x, __nv_bfloat16* __restrict__ xin, int L) {
  const int t = blockIdx.x, c = threadIdx.x;   // 320 threads
  const int part = c / NMEL, ch = c % NMEL;
#pragma unroll
  for (int b = 0; b < 2; ++b) {
    float v = 0.f;
    if (part == 0) { if (in.x[b]) v = in.x[b][(size_t)ch * L + t]; }
    else if (part == 1) { if (in.mu[b]) v = in.mu[b][(size_t)ch * L + t]; }
    else if (part == 2) { if (in.spks[b]) v = in.spks[b][ch]; }
    else { if (in.cond[b]) v = in.cond[b][(size_t)ch * L + t]; }
    if (part == 0 && b == 0 && x) x[(size_t)t * NMEL + ch] = v;
    xin[((size_t)b * L + t) * CIN + c] = __float2bfloat16_rn(v);
  }
}
// est [2][L][80] -> out [2][80][L] (the reference's channel-first layout)
__global__ void k_unpack(const float* __restrict__ est, float* __restrict__ out, int L) {
  const int t = blockIdx.x, ch = threadIdx.x;
#pragma unroll
  for (int b = 0; b < 2; ++b) out[((size_t)b * NMEL + ch) * L + t] = est[((size_t)b * L + t) * NMEL + ch];
}
// v = (1 + r) v_c - r v_u;  x += dt v  (flow_matching.py:115-117)
__global__ void k_euler(const float* __restrict__ est, float* __restrict__ x, __nv_bfloat16* __restrict__ xin,
                        float* __restrict__ mel_out, float dt, float rate, int L, int last) {
  const int t = blockIdx.x, ch = threadIdx.x;   // 80 threads
  const float vc = est[(size_t)t * NMEL + ch], vu = est[((size_t)L + t) * NMEL + ch];
  const float v = (1.0f + rate) * vc - rate * vu;
  const float xn = x[(size_t)t * NMEL + ch] + dt * v;
  x[(size_t)t * NMEL + ch] = xn;
  const __nv_bfloat16 xb = __float2bfloat16_rn(xn);
  xin[(size_t)t * CIN + ch] = xb;
  xin[((size_t)L + t) * CIN + ch] = xb;
  if (last) mel_out[(size_t)ch * L + t] = xn;
}

constexpr int CFM_MAX_GRAPHS = 8;      // cached executable solve graphs (LRU)
constexpr int CFM_MAX_PLANS = 64;      // cached per-length op plans (331 tensor maps each, host memory)
// ---- batched solve (nb utterances, rows 2u = conditional / 2u + 1 = unconditional of utterance u, all padded to L) ----
// z, mu, cond [nb][80][L], spks [nb][80]  ->  x [nb][L][80] fp32 state, xin [2 nb][L][320] bf16 (flow_matching.py:95-108)
__global__ void k_pack_b(const float* __restrict__ z, const float* __restrict__ mu, const float* __restrict__ spks,
                         const float* __restrict__ cond, float* __restrict__ x, __nv_bfloat16* __restrict__ xin, int L) {
  const int t = blockIdx.x, u = blockIdx.y, c = threadIdx.x;   // 320 threads
  const int part = c / NMEL, ch = c % NMEL;
  const size_t o = ((size_t)u * NMEL + ch) * L + t;
  float v;
  if (part == 0) v = z[o];
  else if (part == 1) v = mu[o];
  else if (part == 2) v = spks[(size_t)u * NMEL + ch];
  else v = cond[o];
  if (part == 0) x[((size_t)u * L + t) * NMEL + ch] = v;
  xin[((size_t)(2 * u) * L + t) * CIN + c] = __float2bfloat16_rn(v);
  xin[((size_t)(2 * u + 1) * L + t) * CIN + c] = __float2bfloat16_rn(part == 0 ? v : 0.f);
}
__global__ void k_euler_b(const float* __restrict__ est, float* __restrict__ x, __nv_bfloat16* __restrict__ xin,
                          float* __restrict__ mel_out, float dt, float rate, int L, int last) {
  const int t = blockIdx.x, u = blockIdx.y, ch = threadIdx.x;   // 80 threads
  const float vc = est[((size_t)(2 * u) * L + t) * NMEL + ch], vu = est[((size_t)(2 * u + 1) * L + t) * NMEL + ch];
  const float v = (1.0f + rate) * vc - rate * vu;
  const size_t xo = ((size_t)u * L + t) * NMEL + ch;
  const float xn = x[xo] + dt * v;
  x[xo] = xn;
  const __nv_bfloat16 xb = __float2bfloat16_rn(xn);
  xin[((size_t)(2 * u) * L + t) * CIN + ch] = xb;
  xin[((size_t)(2 * u + 1) * L + t) * CIN + ch] = xb;
  if (last) mel_out[((size_t)u * NMEL + ch) * L + t] = xn;
}

struct Op { int kind; GemmDesc g; int temb_res; FfnDesc f; };   // kind 0 gemm, 1 attention, 2 skip copy, 3 fused feed-forward
struct Plan { int L, nb; std::vector<Op> ops; };

}  // namespace

struct cbx_cfm {
  int device, max_len, max_batch;
  int* lens_dev;         // [2 max_batch] key length of every batch row (padded batches)
  const __nv_bfloat16* wbf; const float* wf;
  std::vector<long long> bo, fo;
  void* slab;
  __nv_bfloat16 *xin, *cur_a, *cur_b, *cat, *ln, *qkv, *att, *ff, *h1;
  float *res, *h2, *est, *x, *mish_temb, *tvec, *tvals, *ffn_red;
  int ffn_red_tiles;
  int max_steps;
  std::map<long long, Plan> plans;      // key = L * 256 + nb
  long long launches;
  // a solve seen twice with the same (L, schedule, cfg rate) is captured into a CUDA graph: ~3 400 launches become
  // one graph launch (the inter-kernel gaps were a fifth of the solve).  Caller buffers are staged so the graph's
  // baked pointers stay valid.
  struct SolveGraph { int L, nb, n, seen; float cfg; std::vector<float> t, dt; cudaGraphExec_t exec; long long launches; };
  std::vector<SolveGraph> graphs;
  float *st_z, *st_mu, *st_spks, *st_cond, *st_out;
  cudaStream_t cap_stream;
  std::vector<float> tv_cache; cudaStream_t tv_stream;     // t values whose time vectors sit in tvec (computed on tv_stream)
};

static const __nv_bfloat16* BFW(cbx_cfm* h, int i) { return h->wbf + h->bo[i]; }
static const float* FW(cbx_cfm* h, int i) { return h->wf + h->fo[i]; }

// ---- the fixed op sequence of one estimator call for mel length L ----
static int build_plan(cbx_cfm* h, int L, int nb, Plan** out) {
  const long long key = (long long)L * 256 + nb;
  auto it = h->plans.find(key);
  if (it != h->plans.end()) { *out = &it->second; return CBX_OK; }
  Plan pl; pl.L = L; pl.nb = nb;
  int rc = CBX_OK;
  // few row tiles (one utterance: 2 x ceil(L / 128)): the feed-forward pair runs as one 8-CTA-cluster kernel per tile;
  // large padded batches keep the two full-width GEMMs (CBX_CFM_FFN=0 / CBX_CFM_FFN_MAXTILES=n for A/B)
  static const int ffn_on = [] { const char* v = getenv("CBX_CFM_FFN"); return v ? atoi(v) : 1; }();
  static const int ffn_max_tiles = [] { const char* v = getenv("CBX_CFM_FFN_MAXTILES"); return v ? atoi(v) : 64; }();
  const bool fused_ffn = ffn_on && 2 * nb * ((L + 127) / 128) <= ffn_max_tiles;
  static const int out_on = [] { const char* v = getenv("CBX_CFM_OUTFFN"); return v ? atoi(v) : 1; }();
  const bool fused_out = out_on != 0;          // the attention out-projection inside the feed-forward kernel (A/B switch)
  auto G = [&](const __nv_bfloat16* A, int lda, int wi, int N, int K, int taps, const GemmEpi& e, int temb_res) {
    Op op; op.kind = 0; op.temb_res = temb_res;
    if (gemm_make(&op.g, A, lda, BFW(h, wi), 2 * nb, L, N, K, taps, e)) rc = CBX_ERR;
    pl.ops.push_back(op);
  };
  auto resnet = [&](int r, const __nv_bfloat16* X, int ldx, int cin, int first_tfm) {
    const int f = F_RES + r * 7;
    GemmEpi e1{}; e1.bias = FW(h, f + 0); e1.ln_g = FW(h, f + 1); e1.ln_b = FW(h, f + 2); e1.out_bf16 = h->h1; e1.ld_bf16 = C;
    G(X, ldx, BF_RES + r * 3 + 0, C, cin, 3, e1, r);
    GemmEpi e2{}; e2.bias = FW(h, f + 3); e2.ln_g = FW(h, f + 4); e2.ln_b = FW(h, f + 5); e2.out_f32 = h->h2; e2.ld_f32 = C;
    G(h->h1, C, BF_RES + r * 3 + 1, C, C, 3, e2, -1);
    GemmEpi e3{}; e3.bias = FW(h, f + 6); e3.resid = h->h2; e3.ld_resid = C; e3.out_f32 = h->res; e3.ld_f32 = C;
    e3.ln2_g = FW(h, F_TFM + first_tfm * 7 + 0); e3.ln2_b = FW(h, F_TFM + first_tfm * 7 + 1); e3.out_ln = h->ln; e3.ld_ln = C;
    G(X, ldx, BF_RES + r * 3 + 2, C, cin, 1, e3, -1);
  };
  auto tfm = [&](int i, bool last, __nv_bfloat16* dst, int ld_dst) {
    const int f = F_TFM + i * 7, w = BF_TFM + i * 4;
    GemmEpi eq{}; eq.out_bf16 = h->qkv; eq.ld_bf16 = 1536;
    G(h->ln, C, w + 0, 1536, C, 1, eq, -1);
    Op at; at.kind = 1; at.temb_res = -1; pl.ops.push_back(at);
    if (!(fused_ffn && fused_out)) {
      GemmEpi eo{}; eo.bias = FW(h, f + 2); eo.resid = h->res; eo.ld_resid = C; eo.out_f32 = h->res; eo.ld_f32 = C;
      eo.ln2_g = FW(h, f + 3); eo.ln2_b = FW(h, f + 4); eo.out_ln = h->ln; eo.ld_ln = C;
      G(h->att, 512, w + 1, C, 512, 1, eo, -1);
    }
    if (fused_ffn) {
      // FF1 + GELU + FF2 + residual (+ the next block's norm1) as ONE cluster kernel (csrc/s3_ffn.cu)
      Op op; op.kind = 3; op.temb_res = -1;
      if (ffn_make(&op.f, h->ln, C, BFW(h, w + 2), BFW(h, w + 3), 2 * nb, L)) rc = CBX_ERR;
      op.f.b1 = FW(h, f + 5); op.f.b2 = FW(h, f + 6);
      op.f.resid = h->res; op.f.ld_resid = C; op.f.out_f32 = h->res; op.f.ld_f32 = C;
      op.f.ln_g = op.f.ln_b = nullptr; op.f.out_ln = nullptr; op.f.ld_ln = C; op.f.out_bf16 = nullptr; op.f.ld_bf16 = 0;
      if (!last) { op.f.ln_g = FW(h, f + 7); op.f.ln_b = FW(h, f + 8); op.f.out_ln = h->ln; }
      else { op.f.out_bf16 = dst; op.f.ld_bf16 = ld_dst; }
      static const int red_l2 = [] { const char* v = getenv("CBX_FFN_RED_L2"); return v ? atoi(v) : 1; }();
      if (red_l2 && 2 * nb * ((L + 127) / 128) <= h->ffn_red_tiles) op.f.red = h->ffn_red;
      if (fused_out) {      // the attention out-projection + residual + norm3 run inside the same cluster kernel
        if (ffn_make_out(&op.f, h->att, 512, BFW(h, w + 1))) rc = CBX_ERR;
        op.f.bo = FW(h, f + 2); op.f.g0 = FW(h, f + 3); op.f.b0 = FW(h, f + 4);
        op.f.resid0 = h->res; op.f.ld_res0 = C; op.f.out0_f32 = h->res; op.f.ld_out0 = C; op.f.xbuf = h->ln; op.f.ld_x = C;
      }
      pl.ops.push_back(op);
      return;
    }
    GemmEpi e1{}; e1.bias = FW(h, f + 5); e1.gelu = 1; e1.out_bf16 = h->ff; e1.ld_bf16 = 1024;
    G(h->ln, C, w + 2, 1024, C, 1, e1, -1);
    GemmEpi e2{}; e2.bias = FW(h, f + 6); e2.resid = h->res; e2.ld_resid = C; e2.out_f32 = h->res; e2.ld_f32 = C;
    if (!last) { e2.ln2_g = FW(h, f + 7); e2.ln2_b = FW(h, f + 8); e2.out_ln = h->ln; e2.ld_ln = C; }   // next block's norm1
    else { e2.out_bf16 = dst; e2.ld_bf16 = ld_dst; }
    G(h->ff, 1024, w + 3, C, 1024, 1, e2, -1);
  };
  auto group = [&](int r, __nv_bfloat16* dst, int ld_dst) {
    for (int j = 0; j < 4; ++j) tfm(r * 4 + j, j == 3, dst, ld_dst);
  };
  // down block
  resnet(0, h->xin, CIN, CIN, 0);
  group(0, h->cur_a, C);
  { Op cp; cp.kind = 2; cp.temb_res = -1; pl.ops.push_back(cp); }          // skip connection -> cat[:, 256:512]
  { GemmEpi e{}; e.bias = FW(h, F_TAIL + 0); e.out_bf16 = h->cur_b; e.ld_bf16 = C; G(h->cur_a, C, BF_TAIL + 0, C, C, 3, e, -1); }
  // mid blocks
  for (int m = 1; m <= 12; ++m) {
    resnet(m, h->cur_b, C, C, m * 4);
    if (m < 12) group(m, h->cur_b, C); else group(m, h->cat, 2 * C);
  }
  // up block: input = [x | skip]
  resnet(13, h->cat, 2 * C, 2 * C, 52);
  group(13, h->cur_a, C);
  { GemmEpi e{}; e.bias = FW(h, F_TAIL + 1); e.out_bf16 = h->cur_b; e.ld_bf16 = C; G(h->cur_a, C, BF_TAIL + 1, C, C, 3, e, -1); }
  { GemmEpi e{}; e.bias = FW(h, F_TAIL + 2); e.ln_g = FW(h, F_TAIL + 3); e.ln_b = FW(h, F_TAIL + 4); e.out_bf16 = h->h1; e.ld_bf16 = C;
    G(h->cur_b, C, BF_TAIL + 2, C, C, 3, e, -1); }
  { GemmEpi e{}; e.bias = FW(h, F_TAIL + 5); e.out_f32 = h->est; e.ld_f32 = NMEL; G(h->h1, C, BF_TAIL + 3, NMEL, C, 1, e, -1); }
  if (rc) return CBX_ERR;
  if ((int)h->plans.size() >= CFM_MAX_PLANS) h->plans.erase(h->plans.begin());   // plans are rebuilt on demand (host work only)
  auto ins = h->plans.emplace(key, std::move(pl));
  *out = &ins.first->second;
  return CBX_OK;
}

static int run_estimator(cbx_cfm* h, Plan* pl, int step, const int* lens, cudaStream_t st) {
  const int L = pl->L, B = 2 * pl->nb;
  for (Op& op : pl->ops) {
    if (op.kind == 0) {
      if (op.temb_res >= 0) {
        GemmDesc d = op.g;
        d.epi.temb = h->tvec + ((size_t)(step * NRES + op.temb_res) * 2) * C;      // [2][C] copies; the GEMM reads the first
        if (gemm_launch(d, st)) return CBX_ERR;
      } else if (gemm_launch(op.g, st)) return CBX_ERR;
    } else if (op.kind == 3) {
      if (ffn_launch(op.f, st)) return CBX_ERR;
    } else if (op.kind == 1) {
      if (s3_attn_launch(h->qkv, h->att, B, L, 1536, lens, st)) return CBX_ERR;
    } else {
      CBX_CUDA(cudaMemcpy2DAsync(h->cat + C, 2 * C * 2, h->cur_a, C * 2, C * 2, (size_t)B * L, cudaMemcpyDeviceToDevice, st));
    }
    h->launches += 1;
  }
  return CBX_OK;
}

// Time embedding MLP + the per-resnet time vectors of every step (decoder.py:256-257, 67-70).  They depend on the t values
// only — the same cosine schedule for every utterance — so the handle keeps the last schedule's vectors: a solve with the
// same t values on the same stream skips the two kernels (0.8 ms of a 30 ms solve: k_temb is a latency-bound 320 -> 1024 ->
// 1024 MLP on n CTAs).
static int time_vectors(cbx_cfm* h, const float* t_host, int n, cudaStream_t st) {
  CBX_CHECK(n >= 1 && n <= h->max_steps, "cfm: n_steps %d out of range (max %d)", n, h->max_steps);
  if (h->tv_stream == st && (int)h->tv_cache.size() == n && std::equal(t_host, t_host + n, h->tv_cache.begin())) return CBX_OK;
  h->tv_cache.assign(t_host, t_host + n); h->tv_stream = st;
  CBX_CUDA(cudaMemcpyAsync(h->tvals, t_host, n * sizeof(float), cudaMemcpyHostToDevice, st));
  k_temb<<<n, 1024, 0, st>>>(h->tvals, FW(h, F_TIME + 0), FW(h, F_TIME + 1), FW(h, F_TIME + 2), FW(h, F_TIME + 3), h->mish_temb);
  k_tvec<<<dim3(NRES, n), 256, 0, st>>>(h->mish_temb, FW(h, F_TIME + 4), FW(h, F_TIME + 5), h->tvec);
  CBX_CUDA(cudaGetLastError());
  h->launches += 2;
  return CBX_OK;
}

// =================================================================================================
// C-ABI (include/cbx.h)
// =================================================================================================
extern "C" int cbx_cfm_create_batched(int device, int max_len, int max_batch, const CbxCfmWeights* w, cbx_cfm** out);
extern "C" int cbx_cfm_create(int device, int max_len, const CbxCfmWeights* w, cbx_cfm** out) {
  return cbx_cfm_create_batched(device, max_len, 1, w, out);
}
extern "C" int cbx_cfm_create_batched(int device, int max_len, int max_batch, const CbxCfmWeights* w, cbx_cfm** out) {
  CBX_CHECK(w && out && w->bf16_blob && w->f32_blob && w->bf16_off && w->f32_off, "cbx_cfm_create: null argument");
  CBX_CHECK(max_len >= 1, "cbx_cfm_create: max_len must be positive");
  CBX_CHECK(max_batch >= 1 && max_batch <= 128, "cbx_cfm_create: max_batch %d out of range (1..128)", max_batch);
  CBX_ON_DEVICE(device);
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, device));
  CBX_CHECK(prop.major == 10, "cbx_cfm_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  cbx_cfm* h = new cbx_cfm();
  h->device = device; h->max_len = max_len; h->max_batch = max_batch; h->launches = 0; h->max_steps = 64;
  h->wbf = static_cast<const __nv_bfloat16*>(w->bf16_blob); h->wf = w->f32_blob;
  h->bo.assign(w->bf16_off, w->bf16_off + BF_COUNT);
  h->fo.assign(w->f32_off, w->f32_off + F_COUNT);
  const size_t R = (size_t)2 * max_len * max_batch, U1 = (size_t)max_len * max_batch;     // batch rows x frames
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 1023) / 1024 * 1024; return o; };
  const size_t o_xin = take(R * CIN * 2), o_ca = take(R * C * 2), o_cb = take(R * C * 2), o_cat = take(R * 2 * C * 2);
  const size_t o_ln = take(R * C * 2), o_qkv = take(R * 1536 * 2), o_att = take(R * 512 * 2), o_ff = take(R * 1024 * 2);
  const size_t o_h1 = take(R * C * 2), o_res = take(R * C * 4), o_h2 = take(R * C * 4), o_est = take(R * NMEL * 4);
  const size_t o_x = take(U1 * NMEL * 4), o_mt = take((size_t)h->max_steps * 1024 * 4);
  const size_t o_tv = take((size_t)h->max_steps * NRES * 2 * C * 4), o_t = take(h->max_steps * 4);
  const size_t o_sz = take(U1 * NMEL * 4), o_sm = take(U1 * NMEL * 4), o_ss = take((size_t)max_batch * NMEL * 4);
  const size_t o_sc = take(U1 * NMEL * 4), o_so = take(U1 * NMEL * 4), o_len = take((size_t)2 * max_batch * 4);
  // fused feed-forward: the 8 partial tiles of a row tile meet in this scratch (1 MB per row tile, L2-resident)
  h->ffn_red_tiles = std::min(64, 2 * max_batch * ((max_len + 127) / 128));
  const size_t o_red = take((size_t)h->ffn_red_tiles * 8 * 128 * 256 * 4);
  if (cudaMalloc(&h->slab, off) != cudaSuccess) {
    cbx_set_error("cbx_cfm_create: cudaMalloc of %zu bytes failed", off);
    delete h;
    return CBX_ERR;
  }
  cudaMemset(h->slab, 0, off);
  uint8_t* s = static_cast<uint8_t*>(h->slab);
  h->xin = (__nv_bfloat16*)(s + o_xin); h->cur_a = (__nv_bfloat16*)(s + o_ca); h->cur_b = (__nv_bfloat16*)(s + o_cb);
  h->cat = (__nv_bfloat16*)(s + o_cat); h->ln = (__nv_bfloat16*)(s + o_ln); h->qkv = (__nv_bfloat16*)(s + o_qkv);
  h->att = (__nv_bfloat16*)(s + o_att); h->ff = (__nv_bfloat16*)(s + o_ff); h->h1 = (__nv_bfloat16*)(s + o_h1);
  h->ffn_red = (float*)(s + o_red);
  h->res = (float*)(s + o_res); h->h2 = (float*)(s + o_h2); h->est = (float*)(s + o_est); h->x = (float*)(s + o_x);
  h->mish_temb = (float*)(s + o_mt); h->tvec = (float*)(s + o_tv); h->tvals = (float*)(s + o_t);
  h->st_z = (float*)(s + o_sz); h->st_mu = (float*)(s + o_sm); h->st_spks = (float*)(s + o_ss); h->st_cond = (float*)(s + o_sc);
  h->st_out = (float*)(s + o_so); h->lens_dev = (int*)(s + o_len);
  h->cap_stream = nullptr;
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_cfm_destroy(cbx_cfm* h) {
  if (!h) return CBX_OK;
  for (auto& g : h->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

extern "C" int cbx_cfm_estimator(cbx_cfm* h, const float* x, const float* mu, const float* spks, const float* cond, float t,
                                 int L, float* out, void* stream) {
  CBX_CHECK(h && x && mu && spks && cond && out, "cbx_cfm_estimator: null argument");
  CBX_CHECK(L >= 1 && L <= h->max_len, "cbx_cfm_estimator: L=%d exceeds max_len %d", L, h->max_len);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->device);
  Plan* pl;
  if (build_plan(h, L, 1, &pl)) return CBX_ERR;
  if (time_vectors(h, &t, 1, st)) return CBX_ERR;
  const size_t n = (size_t)NMEL * L;
  PackIn in{{x, x + n}, {mu, mu + n}, {spks, spks + NMEL}, {cond, cond + n}};
  k_pack<<<L, CIN, 0, st>>>(in, nullptr, h->xin, L);
  if (run_estimator(h, pl, 0, nullptr, st)) return CBX_ERR;
  k_unpack<<<L, NMEL, 0, st>>>(h->est, out, L);
  CBX_CUDA(cudaGetLastError());
  h->launches += 2;
  return CBX_OK;
}

// the kernel sequence of one Euler solve on the handle's own buffers (stream-ordered; capturable)
static int solve_body(cbx_cfm* h, Plan* pl, const float* z, const float* mu, const float* spks, const float* cond, int L,
                      int n_steps, const float* dt_vals, float cfg_rate, float* mel_out, const int* lens, cudaStream_t st) {
  const int nb = pl->nb;
  k_pack_b<<<dim3(L, nb), CIN, 0, st>>>(z, mu, spks, cond, h->x, h->xin, L);
  h->launches += 1;
  for (int s = 0; s < n_steps; ++s) {
    if (run_estimator(h, pl, s, lens, st)) return CBX_ERR;
    k_euler_b<<<dim3(L, nb), NMEL, 0, st>>>(h->est, h->x, h->xin, mel_out, dt_vals[s], cfg_rate, L, s == n_steps - 1);
    h->launches += 1;
  }
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

extern "C" int cbx_cfm_solve_batch(cbx_cfm* h, int nb, const float* z, const float* mu, const float* spks, const float* cond,
                                   const int32_t* lens, int L, int n_steps, const float* t_vals, const float* dt_vals,
                                   float cfg_rate, float* mel_out, void* stream);

extern "C" int cbx_cfm_solve(cbx_cfm* h, const float* z, const float* mu, const float* spks, const float* cond, int L,
                             int n_steps, const float* t_vals, const float* dt_vals, float cfg_rate, float* mel_out,
                             void* stream) {
  return cbx_cfm_solve_batch(h, 1, z, mu, spks, cond, nullptr, L, n_steps, t_vals, dt_vals, cfg_rate, mel_out, stream);
}

extern "C" int cbx_cfm_solve_batch(cbx_cfm* h, int nb, const float* z, const float* mu, const float* spks, const float* cond,
                                   const int32_t* lens, int L, int n_steps, const float* t_vals, const float* dt_vals,
                                   float cfg_rate, float* mel_out, void* stream) {
  CBX_CHECK(h && z && mu && spks && cond && t_vals && dt_vals && mel_out, "cbx_cfm_solve: null argument");
  CBX_CHECK(nb >= 1 && nb <= h->max_batch, "cbx_cfm_solve: batch %d exceeds max_batch %d", nb, h->max_batch);
  CBX_CHECK(L >= 1 && L <= h->max_len, "cbx_cfm_solve: L=%d exceeds max_len %d", L, h->max_len);
  CBX_CHECK(n_steps >= 1 && n_steps <= h->max_steps, "cfm: n_steps %d out of range (max %d)", n_steps, h->max_steps);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->device);
  Plan* pl;
  if (build_plan(h, L, nb, &pl)) return CBX_ERR;
  if (time_vectors(h, t_vals, n_steps, st)) return CBX_ERR;      // outside the captured graph: usually a no-op (cached schedule)
  // key lengths per batch row (rows 2u, 2u + 1 = utterance u); NULL = every utterance is L frames long
  const int* lens_dev = nullptr;
  if (lens) {
    std::vector<int> rows(2 * nb);
    for (int u = 0; u < nb; ++u) {
      CBX_CHECK(lens[u] >= 1 && lens[u] <= L, "cbx_cfm_solve_batch: lens[%d] = %d outside [1, L = %d]", u, lens[u], L);
      rows[2 * u] = rows[2 * u + 1] = lens[u];
    }
    CBX_CUDA(cudaMemcpyAsync(h->lens_dev, rows.data(), rows.size() * sizeof(int), cudaMemcpyHostToDevice, st));   // pageable: staged before return
    lens_dev = h->lens_dev;
  }
  static const bool use_graph = [] { const char* v = getenv("CBX_CFM_GRAPH"); return !(v && v[0] == '0'); }();
  cbx_cfm::SolveGraph* g = nullptr;
  const int masked = lens ? 1 : 0;         // the graph bakes lens_dev (or NULL) into the attention launches
  if (use_graph) {
    for (auto& c : h->graphs)
      if (c.L == L && c.nb == nb * 2 + masked && c.n == n_steps && c.cfg == cfg_rate && c.t == std::vector<float>(t_vals, t_vals + n_steps) &&
          c.dt == std::vector<float>(dt_vals, dt_vals + n_steps)) { g = &c; break; }
    if (!g) {
      // bounded: a long-lived server sees ~1 250 distinct mel lengths; the oldest entry (and its ~3 400-node
      // executable graph) is dropped, its length simply runs eagerly again until it is seen twice
      if ((int)h->graphs.size() >= CFM_MAX_GRAPHS) {
        if (h->graphs.front().exec) cudaGraphExecDestroy(h->graphs.front().exec);
        h->graphs.erase(h->graphs.begin());
      }
      h->graphs.push_back({L, nb * 2 + masked, n_steps, 0, cfg_rate, std::vector<float>(t_vals, t_vals + n_steps),
                           std::vector<float>(dt_vals, dt_vals + n_steps), nullptr, 0});
      g = &h->graphs.back();
    } else if (g != &h->graphs.back()) {          // most recently used last
      cbx_cfm::SolveGraph tmp = std::move(*g);
      h->graphs.erase(h->graphs.begin() + (g - &h->graphs[0]));
      h->graphs.push_back(std::move(tmp));
      g = &h->graphs.back();
    }
    g->seen += 1;
  }
  if (!g || g->seen < 2) return solve_body(h, pl, z, mu, spks, cond, L, n_steps, dt_vals, cfg_rate, mel_out, lens_dev, st);   // eager
  const size_t nb_bytes = (size_t)nb * NMEL * L * sizeof(float);
  if (!g->exec) {
    if (!h->cap_stream) CBX_CUDA(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
    const long long before = h->launches;
    cudaGraph_t graph = nullptr;
    CBX_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
    const int rc = solve_body(h, pl, h->st_z, h->st_mu, h->st_spks, h->st_cond, L, n_steps, dt_vals, cfg_rate, h->st_out, lens_dev, h->cap_stream);
    const cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
    g->launches = h->launches - before;
    h->launches = before;
    if (rc || ce != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      cbx_set_error("cbx_cfm_solve: graph capture failed (%s)", cudaGetErrorString(ce));
      return CBX_ERR;
    }
    const cudaError_t ie = cudaGraphInstantiate(&g->exec, graph, 0);
    cudaGraphDestroy(graph);
    CBX_CHECK(ie == cudaSuccess, "cbx_cfm_solve: cudaGraphInstantiate failed (%s)", cudaGetErrorString(ie));
  }
  CBX_CUDA(cudaMemcpyAsync(h->st_z, z, nb_bytes, cudaMemcpyDeviceToDevice, st));
  CBX_CUDA(cudaMemcpyAsync(h->st_mu, mu, nb_bytes, cudaMemcpyDeviceToDevice, st));
  CBX_CUDA(cudaMemcpyAsync(h->st_spks, spks, (size_t)nb * NMEL * sizeof(float), cudaMemcpyDeviceToDevice, st));
  CBX_CUDA(cudaMemcpyAsync(h->st_cond, cond, nb_bytes, cudaMemcpyDeviceToDevice, st));
  CBX_CUDA(cudaGraphLaunch(g->exec, st));
  CBX_CUDA(cudaMemcpyAsync(mel_out, h->st_out, nb_bytes, cudaMemcpyDeviceToDevice, st));
  h->launches += g->launches;
  return CBX_OK;
}

namespace cbx { void gemm_set_debug(unsigned long long* buf); void ffn_set_debug(unsigned long long* buf); }
/* Debug aid (not in cbx.h): every following GEMM launch writes 8 %globaltimer stamps of CTA (0,0) to buf and advances it. */
extern "C" int cbx_dbg_gemm_stamps(unsigned long long* buf) { cbx::gemm_set_debug(buf); return 0; }
namespace cbx { int s3_attn_tc_launch(const __nv_bfloat16*, __nv_bfloat16*, int, int, int, const int*, cudaStream_t); }
/* Debug aid (not in cbx.h): the estimator's attention kernel alone — qkv bf16 [B][L][1536] -> out bf16 [B][L][512];
   tc != 0: the tcgen05 kernel, 0: the mma.sync kernel (tests/test_s3_attn_gpu.py). */
// test hook: y[i] = gelu_f(x[i]), the GELU every estimator epilogue uses (tests/test_cfm_gpu.py pins its error against erf)
__global__ void k_dbg_gelu(const float* __restrict__ x, float* __restrict__ y, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = cbx::gelu_f(x[i]);
}
extern "C" int cbx_dbg_gelu(const float* x, float* y, int n, void* stream) {
  CBX_CHECK(x && y && n >= 1, "cbx_dbg_gelu: bad argument");
  k_dbg_gelu<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, n);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

extern "C" int cbx_dbg_s3_attn(const void* qkv, void* out, int B, int L, const int* lens, int tc, void* stream) {
  CBX_CHECK(qkv && out && B >= 1 && L >= 1, "cbx_dbg_s3_attn: bad argument");
  auto q = static_cast<const __nv_bfloat16*>(qkv); auto o = static_cast<__nv_bfloat16*>(out);
  if (tc) return cbx::s3_attn_tc_launch(q, o, B, L, 1536, lens, static_cast<cudaStream_t>(stream));
  return cbx::s3_attn_launch_mma(q, o, B, L, 1536, lens, static_cast<cudaStream_t>(stream));
}
namespace cbx { void attn_tc_set_debug(unsigned long long* buf); }
extern "C" int cbx_dbg_attn_stamps(unsigned long long* buf) { cbx::attn_tc_set_debug(buf); return 0; }
extern "C" int cbx_dbg_ffn_stamps(unsigned long long* buf) { cbx::ffn_set_debug(buf); return 0; }

extern "C" int cbx_cfm_cache_sizes(cbx_cfm* h, int32_t* n_graphs, int32_t* n_plans) {
  CBX_CHECK(h, "cbx_cfm_cache_sizes: null handle");
  if (n_graphs) *n_graphs = (int32_t)h->graphs.size();
  if (n_plans) *n_plans = (int32_t)h->plans.size();
  return CBX_OK;
}

extern "C" int cbx_cfm_launches(cbx_cfm* h, int64_t* launches) {
  CBX_CHECK(h && launches, "cbx_cfm_launches: null argument");
  *launches = h->launches;
  return CBX_OK;
}
