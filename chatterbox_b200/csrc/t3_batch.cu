// Batched T3 decode engine for sm_100a: U concurrent utterances ("slots"), every decode step one pass of ALL slots'
// CFG pairs (M = 2 U rows <= 128) through each weight matrix on the tcgen05 GEMM path, replayed as ONE CUDA graph.
//
// What it replaces: the reference serves one request at a time — src/server/server.py:189-299 calls
// `tts_model.generate` synchronously per request (:223,232) and swaps the global `model.conds`
// (src/server/voice_manager.py:370-379) — so every request pays the 1.02 GB weight stream per token on its own
// (src/chatterbox/models/t3/t3.py:319-367).  BASELINE config 4 ("64 concurrent utterances per GPU") instead shares
// the weight stream between the slots of a step; the per-utterance state (KV slab, generated-id set, Philox stream,
// sampling knobs, EOS flag) lives in per-slot arrays on the device.
//
// One step (= graph replay), all kernels on one stream with programmatic dependent launch:
//   k_b_sample      per slot: CFG, temperature, repetition penalty, min-p, top-p, multinomial / argmax, EOS, next-token
//                   embedding -> fp32 residual rows (t3.py:320-356; same arithmetic as sample_phase of t3_decode.cu)
//   per layer       k_rms_hilo -> q|k|v GEMM -> k_b_attn (RoPE, KV append, split-KV attention over the slot's own slab
//                   [+ k_b_attn_combine]) -> o_proj GEMM + residual -> k_rms_hilo -> gate/up GEMM + SwiGLU -> down GEMM
//   final           k_rms_hilo -> speech_head GEMM -> logits [2 U][8208]
// Activations travel as hi / lo bf16 planes (~16 mantissa bits) like the prompt prefill (t3_prefill.cu), weights are
// the plain bf16 matrices of CbxT3Weights.wplain.  The step is HBM-bound: weights once (1.02 GB) + every running slot's
// KV (245 760 B x context) per step.
#include "t3_prefill.h"
#include "s3_gemm.cuh"
#include "t3_rows.cuh"
#include "../../include/cbx.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

namespace cbx {
namespace {

using namespace t3rows;

constexpr int VOCAB_LD = 8208;              // logits row pitch (speech_head rows padded to a multiple of 16)
constexpr int SEEN_WORDS = (VOCAB + 31) / 32;
constexpr int STOP_SPEECH = 6562, START_SPEECH = 6561;
constexpr int NT = 256;                     // threads of the sampling kernel
constexpr int MAX_SPLITS = 8;               // KV splits of the decode attention at small batch
constexpr int PART = 68;                    // floats of one attention partial: 64 o + m + l (+ pad)

struct SlotState {     // one per slot, device memory
  int pos;             // KV cache length (positions appended so far)
  int n_gen;           // speech tokens emitted
  int done;            // 1: EOS seen or max_new reached (or never started)
  int fwd;             // 1: this step's forward pass runs for the slot (its rows are live)
  int active;          // 1: a request occupies the slot
  int max_new;
  int has_forced;
  int err;
};

struct BatchParams {
  int U, cap;                        // slots, token capacity per slot
  SlotState* st;
  const CbxSampleParams* sp;         // [U]
  unsigned* seen;                    // [U][SEEN_WORDS]
  const float* logits;               // [2 U][VOCAB_LD]
  const int* mask;                   // [U] slots this call advances
  const int* forced;                 // [U][cap]
  int* tokens;                       // [U][cap]
  float* logits_out;                 // optional [U][cap][2][VOCAB]
  float* x;                          // [2 U][1024] fp32 residual stream
  const float *speech_emb, *speech_pos_emb;
};

__device__ __forceinline__ float philox_uniform(unsigned long long seed, unsigned ctr) {
  unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
  unsigned c0 = ctr, c1 = 0, c2 = 0x9E3779B9u, c3 = 0;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return (c0 >> 8) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float block_sum(float v, float* rs, int warp, int lane) {
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) rs[warp] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) t += rs[w];
  return t;
}

// ---- per-slot logit processing + sampling + next-token embedding (t3.py:320-356) --------------------------------
// Mirrors sample_phase of t3_decode.cu statement by statement (same reduction shapes, same Philox counter), so a slot
// fed the same logits draws the same token as the single-utterance megakernel.
__global__ void __launch_bounds__(NT) k_b_sample(const BatchParams p) {
  __shared__ float sl[VOCAB_LD];
  __shared__ float rs[32];
  __shared__ int ri[32];
  __shared__ unsigned seen_s[SEEN_WORDS];
  const int u = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_launch_dependents();
  pdl_wait();
  SlotState* st = p.st + u;
  float4* xr0 = reinterpret_cast<float4*>(p.x + (size_t)(2 * u) * H);
  float4* xr1 = reinterpret_cast<float4*>(p.x + (size_t)(2 * u + 1) * H);
  const int was_fwd = st->fwd, pos = st->pos + (was_fwd ? 1 : 0), n_gen = st->n_gen;
  const bool running = st->active && !st->done && p.mask[u];
  __syncthreads();                       // every thread has read the slot state before thread 0 rewrites it
  if (!running) {
    if (tid == 0) { st->pos = pos; st->fwd = 0; }
    xr0[tid] = make_float4(0.f, 0.f, 0.f, 0.f);
    xr1[tid] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const CbxSampleParams sp = p.sp[u];
  const bool cfg = sp.cfg_weight > 0.f;
  for (int i = tid; i < SEEN_WORDS; i += NT) seen_s[i] = p.seen[(size_t)u * SEEN_WORDS + i];
  __syncthreads();
  const float* lc = p.logits + (size_t)(2 * u) * VOCAB_LD;
  const float* lu = lc + VOCAB_LD;
  float* lout = p.logits_out ? p.logits_out + ((size_t)u * p.cap + n_gen) * 2 * VOCAB : nullptr;
  // 1. CFG, temperature, repetition penalty; running max (first index on ties)
  float lmax = -INFINITY; int imax = 0x7fffffff;
  for (int i = tid; i < VOCAB; i += NT) {
    const float c = lc[i], uu = lu[i];
    if (lout) { lout[i] = c; lout[VOCAB + i] = uu; }
    float x = cfg ? c + sp.cfg_weight * (c - uu) : c;
    if (sp.temperature != 1.0f) x = x / sp.temperature;
    if ((seen_s[i >> 5] >> (i & 31)) & 1u) x = (x < 0.f) ? x * sp.repetition_penalty : x / sp.repetition_penalty;
    sl[i] = x;
    if (x > lmax) { lmax = x; imax = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float m2 = __shfl_xor_sync(0xffffffffu, lmax, o);
    int i2 = __shfl_xor_sync(0xffffffffu, imax, o);
    if (m2 > lmax || (m2 == lmax && i2 < imax)) { lmax = m2; imax = i2; }
  }
  __syncthreads();
  if (lane == 0) { rs[warp] = lmax; ri[warp] = imax; }
  __syncthreads();
  float M = rs[0]; int iM = ri[0];
#pragma unroll
  for (int w = 1; w < NT / 32; ++w)
    if (rs[w] > M || (rs[w] == M && ri[w] < iM)) { M = rs[w]; iM = ri[w]; }
  int tok = iM;
  if (!sp.greedy && !st->has_forced) {
    // 2. min-p (HF MinPLogitsWarper): drop p_i < min_p * p_max; top-1 kept
    float zloc = 0.f;
    for (int i = tid; i < VOCAB; i += NT) {
      float e = __expf(sl[i] - M);
      if (sp.min_p > 0.f && e < sp.min_p && i != iM) { sl[i] = -INFINITY; e = 0.f; }
      zloc += e;
    }
    float Z = block_sum(zloc, rs, warp, lane);
    // 3. top-p (HF TopPLogitsWarper): drop tokens whose ascending cumulative probability <= 1 - top_p
    if (sp.top_p < 1.0f) {
      const float thr = (1.0f - sp.top_p) * Z;
      unsigned lo = 0u, hi = f2ord(M);
      for (int itb = 0; itb < 32 && lo < hi; ++itb) {
        const unsigned mid = lo + (hi - lo + 1) / 2;
        float s = 0.f;
        for (int i = tid; i < VOCAB; i += NT) {
          const float x = sl[i];
          if (x != -INFINITY && f2ord(x) <= mid) s += __expf(x - M);
        }
        s = block_sum(s, rs, warp, lane);
        if (s <= thr) lo = mid; else hi = mid - 1;
      }
      zloc = 0.f;
      for (int i = tid; i < VOCAB; i += NT) {
        const float x = sl[i];
        if (x != -INFINITY && f2ord(x) <= lo && i != iM) sl[i] = -INFINITY;
        else if (x != -INFINITY) zloc += __expf(x - M);
      }
      Z = block_sum(zloc, rs, warp, lane);
    }
    // 4. multinomial: inverse CDF in index order (contiguous chunk per thread)
    const float target = philox_uniform(sp.seed, (unsigned)n_gen) * Z;
    constexpr int CH = (VOCAB + NT - 1) / NT;
    const int i0 = tid * CH, i1 = min(VOCAB, i0 + CH);
    float loc = 0.f;
    for (int i = i0; i < i1; ++i) loc += (sl[i] == -INFINITY) ? 0.f : __expf(sl[i] - M);
    float inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { float t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    __syncthreads();
    if (lane == 31) rs[warp] = inc;
    if (tid == 0) ri[16] = iM;
    __syncthreads();
    float woff = 0.f;
    for (int w = 0; w < warp; ++w) woff += rs[w];
    const float excl = woff + inc - loc;
    if (target >= excl && target < excl + loc) {
      float acc = excl; int pick = -1;
      for (int i = i0; i < i1; ++i) {
        if (sl[i] == -INFINITY) continue;
        acc += __expf(sl[i] - M);
        pick = i;
        if (target < acc) break;
      }
      if (pick >= 0) ri[16] = pick;
    }
    __syncthreads();
    tok = ri[16];
  }
  if (st->has_forced) tok = p.forced[(size_t)u * p.cap + n_gen];
  const bool stop = sp.stop_on_eos && tok == STOP_SPEECH;
  const bool done = stop || n_gen + 1 >= st->max_new;
  __syncthreads();                       // all reads of has_forced / max_new are done
  if (tid == 0) {
    p.tokens[(size_t)u * p.cap + n_gen] = tok;
    p.seen[(size_t)u * SEEN_WORDS + (tok >> 5)] = seen_s[tok >> 5] | (1u << (tok & 31));
    st->pos = pos; st->n_gen = n_gen + 1; st->done = done ? 1 : 0; st->fwd = done ? 0 : 1;
  }
  // 5. next-token embedding for both CFG rows (t3.py:351-356); a finished slot's rows go inert
  if (done) {
    xr0[tid] = make_float4(0.f, 0.f, 0.f, 0.f);
    xr1[tid] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
    const float4 e = __ldg(reinterpret_cast<const float4*>(p.speech_emb + (size_t)tok * H) + tid);
    const float4 q = __ldg(reinterpret_cast<const float4*>(p.speech_pos_emb + (size_t)(n_gen + 1) * H) + tid);
    const float4 v = make_float4(e.x + q.x, e.y + q.y, e.z + q.z, e.w + q.w);
    xr0[tid] = v;
    xr1[tid] = v;
  }
}

// ---- decode attention over each slot's own KV slab ----------------------------------------------------------------
// grid (16 heads, 2 U rows, S splits), 128 threads.  Row r = 2 slot + CFG row.  RoPE on q / k (rotate_half pairing
// d <-> d ^ 32, llama3-scaled tables), K / V rounded to the slab type and appended at the slot's position by the last
// split, then softmax(q k^T / 8) v over the slot's positions [lo, hi) of this split (+ the new position itself in the
// last split): a warp covers 4 positions per load instruction (8 lanes x 16 bytes = one 64-wide bf16 row), four
// positions per lane group are in flight before the first use, online softmax per lane group, groups and warps merged
// at the end.  S == 1 writes the normalised head output as hi / lo planes; S > 1 writes {o, m, l} partials.
// (reference: HF LlamaAttention eager path, fp32 softmax — t3_hf_backend.py:73-116 with output_attentions=True)
struct AttnParams {
  const float* qkv;                 // [2 U][3072] fp32
  const SlotState* st;
  const float *rope_cos, *rope_sin;
  void *kcache, *vcache;            // [U][layers][2][16][max_pos][64]
  long long slot_stride;            // elements between slots
  int layer, max_pos;
  __nv_bfloat16* out; long long plane;   // hi / lo planes [2][2 U][1024]
  float* part;                      // [2 U][16][S][PART]
};

template <typename KVT> struct KVL;
template <> struct KVL<__nv_bfloat16> {
  struct V { uint4 a; };
  static __device__ __forceinline__ V ld(const __nv_bfloat16* p) { V v; v.a = __ldcs(reinterpret_cast<const uint4*>(p)); return v; }
  static __device__ __forceinline__ void cvt(const V& v, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v.a);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
};
template <> struct KVL<float> {
  struct V { float4 a, b; };
  static __device__ __forceinline__ V ld(const float* p) {
    V v; v.a = __ldcs(reinterpret_cast<const float4*>(p)); v.b = __ldcs(reinterpret_cast<const float4*>(p) + 1); return v;
  }
  static __device__ __forceinline__ void cvt(const V& v, float (&f)[8]) {
    f[0] = v.a.x; f[1] = v.a.y; f[2] = v.a.z; f[3] = v.a.w; f[4] = v.b.x; f[5] = v.b.y; f[6] = v.b.z; f[7] = v.b.w;
  }
};

template <typename KVT>
__global__ void __launch_bounds__(128) k_b_attn(const AttnParams p) {
  __shared__ float q_s[64], k_new[64], v_new[64];
  __shared__ float wpart[4][PART];
  const int head = blockIdx.x, r = blockIdx.y, split = blockIdx.z, S = gridDim.z;
  const int u = r >> 1, b = r & 1, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int fwd = p.st[u].fwd, pos0 = p.st[u].pos;
  float* part = p.part + (((size_t)r * NH + head) * S + split) * PART;
  if (!fwd) {                             // inert rows: the slot is free or finished
    if (S == 1) {
      if (tid < 64) {
        __nv_bfloat16* d = p.out + (size_t)r * H + head * 64 + tid;
        d[0] = __float2bfloat16_rn(0.f); d[p.plane] = __float2bfloat16_rn(0.f);
      }
    } else if (tid < 66) part[tid] = tid == 64 ? -INFINITY : 0.f;
    return;
  }
  KVT* kc = reinterpret_cast<KVT*>(p.kcache) + (size_t)u * p.slot_stride + ((size_t)(p.layer * 2 + b) * NH + head) * p.max_pos * HD;
  KVT* vc = reinterpret_cast<KVT*>(p.vcache) + (size_t)u * p.slot_stride + ((size_t)(p.layer * 2 + b) * NH + head) * p.max_pos * HD;
  const bool last = split == S - 1;
  if (tid < 64) {
    const int d = tid;
    const float* row = p.qkv + (size_t)r * 3072 + head * 64;
    const float q = row[d], qo = row[d ^ 32], k = row[1024 + d], ko = row[1024 + (d ^ 32)], v = row[2048 + d];
    const float cs = __ldg(p.rope_cos + (size_t)pos0 * 32 + (d & 31)), sn = __ldg(p.rope_sin + (size_t)pos0 * 32 + (d & 31));
    const float qr = q * cs + (d < 32 ? -qo : qo) * sn;
    const float kr = KVS<KVT>::round(k * cs + (d < 32 ? -ko : ko) * sn);
    const float vr = KVS<KVT>::round(v);
    q_s[d] = qr * 0.125f; k_new[d] = kr; v_new[d] = vr;
    if (last) {
      KVS<KVT>::store(kc + (size_t)pos0 * HD + d, kr);
      KVS<KVT>::store(vc + (size_t)pos0 * HD + d, vr);
    }
  }
  __syncthreads();
  const int lg = lane >> 3, sub = lane & 7;
  float qv[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) qv[e] = q_s[sub * 8 + e];
  const int chunk = (pos0 + S - 1) / S;
  const int lo = split * chunk, hi = min(pos0, lo + chunk);
  float m = -INFINITY, l = 0.f, o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = 0.f;
  constexpr int UNR = 4;                  // positions per lane group in flight
  for (int base = lo + warp * 4; base < hi; base += 16 * UNR) {     // warp-uniform trip count (full-mask shuffles inside)
    const int i0 = base + lg;
    typename KVL<KVT>::V kr[UNR], vr[UNR];
#pragma unroll
    for (int j = 0; j < UNR; ++j) {
      const int i = min(i0 + 16 * j, hi - 1);            // clamped: a tail slot re-reads the last row, masked below
      kr[j] = KVL<KVT>::ld(kc + (size_t)i * HD + sub * 8);
      vr[j] = KVL<KVT>::ld(vc + (size_t)i * HD + sub * 8);
    }
    float s[UNR];
#pragma unroll
    for (int j = 0; j < UNR; ++j) {
      float kf[8];
      KVL<KVT>::cvt(kr[j], kf);
      float a = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) a += qv[e] * kf[e];
      a += __shfl_xor_sync(0xffffffffu, a, 1);
      a += __shfl_xor_sync(0xffffffffu, a, 2);
      a += __shfl_xor_sync(0xffffffffu, a, 4);
      s[j] = (i0 + 16 * j < hi) ? a : -INFINITY;
    }
    float mn = m;
#pragma unroll
    for (int j = 0; j < UNR; ++j) mn = fmaxf(mn, s[j]);
    const float corr = (mn == -INFINITY) ? 1.f : __expf(m - mn);     // exp(-inf) = 0 on first use; all-masked group: no-op
    l *= corr;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] *= corr;
#pragma unroll
    for (int j = 0; j < UNR; ++j) {
      const float pe = (s[j] == -INFINITY) ? 0.f : __expf(s[j] - mn);
      float vf[8];
      KVL<KVT>::cvt(vr[j], vf);
      l += pe;
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] += pe * vf[e];
    }
    m = mn;
  }
  // the position itself: last split, warp 0, lane group 0
  if (last && warp == 0 && lg == 0) {
    float a = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) a += qv[e] * k_new[sub * 8 + e];
    a += __shfl_xor_sync(0x000000ffu, a, 1);
    a += __shfl_xor_sync(0x000000ffu, a, 2);
    a += __shfl_xor_sync(0x000000ffu, a, 4);
    const float mn = fmaxf(m, a), corr = __expf(m - mn), pe = __expf(a - mn);
    l = l * corr + pe;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = o[e] * corr + pe * v_new[sub * 8 + e];
    m = mn;
  }
  // merge the 4 lane groups (xor 8, 16), then the warps through shared memory
#pragma unroll
  for (int msk = 8; msk <= 16; msk <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, msk), l2 = __shfl_xor_sync(0xffffffffu, l, msk);
    const float mn = fmaxf(m, m2);
    const float c1 = (m == -INFINITY) ? 0.f : __expf(m - mn), c2 = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
    l = l * c1 + l2 * c2;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float o2 = __shfl_xor_sync(0xffffffffu, o[e], msk);
      o[e] = o[e] * c1 + o2 * c2;
    }
    m = mn;
  }
  if (lg == 0) {
#pragma unroll
    for (int e = 0; e < 8; ++e) wpart[warp][sub * 8 + e] = o[e];
    if (sub == 0) { wpart[warp][64] = m; wpart[warp][65] = l; }
  }
  __syncthreads();
  if (tid < 64) {
    const int d = tid;
    float Mx = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) Mx = fmaxf(Mx, wpart[w][64]);
    float Ls = 0.f, O = 0.f;
    if (Mx != -INFINITY) {
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float cw = (wpart[w][64] == -INFINITY) ? 0.f : __expf(wpart[w][64] - Mx);
        Ls += cw * wpart[w][65];
        O += cw * wpart[w][d];
      }
    }
    if (S == 1) {
      const float v = O / Ls;               // the position itself is always present: Ls > 0
      const __nv_bfloat16 hb = __float2bfloat16_rn(v);
      __nv_bfloat16* dst = p.out + (size_t)r * H + head * 64 + d;
      dst[0] = hb;
      dst[p.plane] = __float2bfloat16_rn(v - __bfloat162float(hb));
    } else {
      part[d] = O;
      if (d == 0) { part[64] = Mx; part[65] = Ls; }
    }
  }
}

// grid (16, 2 U), 64 threads: merge the S partials of a (row, head) -> hi / lo planes
__global__ void __launch_bounds__(64) k_b_attn_combine(const float* __restrict__ part, int S, __nv_bfloat16* __restrict__ out,
                                                       long long plane) {
  const int head = blockIdx.x, r = blockIdx.y, d = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const float* pp = part + ((size_t)r * NH + head) * S * PART;
  float Mx = -INFINITY;
  for (int s = 0; s < S; ++s) Mx = fmaxf(Mx, pp[s * PART + 64]);
  float Ls = 0.f, O = 0.f;
  if (Mx != -INFINITY)
    for (int s = 0; s < S; ++s) {
      const float ms = pp[s * PART + 64];
      const float cw = (ms == -INFINITY) ? 0.f : __expf(ms - Mx);
      Ls += cw * pp[s * PART + 65];
      O += cw * pp[s * PART + d];
    }
  const float v = Ls > 0.f ? O / Ls : 0.f;
  const __nv_bfloat16 hb = __float2bfloat16_rn(v);
  __nv_bfloat16* dst = out + (size_t)r * H + head * 64 + d;
  dst[0] = hb;
  dst[plane] = __float2bfloat16_rn(v - __bfloat162float(hb));
}

// slot reset after a prefill: cache length P, nothing generated, generated-id set = {BOS} (t3.py:298)
__global__ void k_b_slot_init(SlotState* st, unsigned* seen, int P, int max_new, int has_forced) {
  const int tid = threadIdx.x;
  for (int i = tid; i < SEEN_WORDS; i += blockDim.x) seen[i] = (i == (START_SPEECH >> 5)) ? (1u << (START_SPEECH & 31)) : 0u;
  if (tid == 0) {
    st->pos = P; st->n_gen = 0; st->done = 0; st->fwd = 0; st->active = 1; st->max_new = max_new; st->has_forced = has_forced;
    st->err = 0;
  }
}
__global__ void k_b_slot_release(SlotState* st) { st->active = 0; st->done = 1; st->fwd = 0; }

// launch with the programmatic-dependent-launch attribute (the kernels above call pdl_wait before their first read)
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

}  // namespace
}  // namespace cbx

// =================================================================================================
// host side of the C-ABI (include/cbx.h, section 1b)
// =================================================================================================
using namespace cbx;

struct cbx_t3b {
  CbxT3BatchConfig cfg;
  CbxT3Weights w;
  int U, M, S;                       // slots, rows (2 U), KV splits of the attention kernel
  void* slab;
  // device buffers (all inside slab)
  void *kc, *vc;
  long long slot_stride;             // KV elements per slot
  float *x, *qkv, *logits, *part, *logits_out;
  __nv_bfloat16 *hl, *act;
  SlotState* st;
  CbxSampleParams* sp;
  unsigned* seen;
  int *forced, *tokens, *step_mask, *pf_state;
  void* prefill_ws;
  long long pf_rows;      // activation rows the prefill workspace holds
  std::vector<GemmDesc> gemms;       // per layer: qkv, o, gate/up, down; then head
  cudaGraphExec_t step_graph;
  cudaStream_t cap_stream;
  long long launches, launches_per_step;
  std::vector<SlotState> host_st;
};

namespace {

struct TileCfg { int qkv_t, qkv_s, o_t, o_s, gu_s, dn_t, dn_s, hd_t, hd_s; };
TileCfg tile_cfg() {
  // column tile / split-K cluster of the five decode GEMMs: every GEMM should put ~100+ CTAs on the 148 SMs, each
  // streaming its own panel of the weight matrix.  CBX_T3B_TILES="qkv_t,qkv_s,o_t,o_s,gu_s,dn_t,dn_s,hd_t,hd_s" overrides.
  TileCfg c{64, 2, 32, 4, 4, 32, 4, 144, 2};
  if (const char* v = getenv("CBX_T3B_TILES")) {
    int a[9];
    if (sscanf(v, "%d,%d,%d,%d,%d,%d,%d,%d,%d", a, a + 1, a + 2, a + 3, a + 4, a + 5, a + 6, a + 7, a + 8) == 9)
      c = TileCfg{a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8]};
  }
  return c;
}

template <typename KVT>
int launch_attn(cbx_t3b* h, int layer, cudaStream_t st) {
  AttnParams a;
  a.qkv = h->qkv; a.st = h->st; a.rope_cos = h->w.rope_cos; a.rope_sin = h->w.rope_sin;
  a.kcache = h->kc; a.vcache = h->vc; a.slot_stride = h->slot_stride; a.layer = layer; a.max_pos = h->cfg.max_pos;
  a.out = h->hl; a.plane = (long long)h->M * H; a.part = h->part;
  CBX_CUDA(launch_pdl(k_b_attn<KVT>, dim3(NH, h->M, h->S), dim3(128), st, a));
  if (h->S > 1)
    CBX_CUDA(launch_pdl(k_b_attn_combine, dim3(NH, h->M), dim3(64), st, (const float*)h->part, h->S, h->hl, (long long)h->M * H));
  return CBX_OK;
}

// the kernel sequence of one decode step (stream-ordered; capturable)
int step_body(cbx_t3b* h, cudaStream_t st) {
  const int M = h->M, L = h->cfg.n_layers;
  const long long hl_plane = (long long)M * H;
  BatchParams bp;
  bp.U = h->U; bp.cap = h->cfg.max_new_cap; bp.st = h->st; bp.sp = h->sp; bp.seen = h->seen; bp.logits = h->logits;
  bp.mask = h->step_mask; bp.forced = h->forced; bp.tokens = h->tokens; bp.logits_out = h->logits_out; bp.x = h->x;
  bp.speech_emb = h->w.speech_emb; bp.speech_pos_emb = h->w.speech_pos_emb;
  CBX_CUDA(launch_pdl(k_b_sample, dim3(h->U), dim3(NT), st, bp));
  long long n = 1;
  for (int l = 0; l < L; ++l) {
    CBX_CUDA(launch_pdl(k_rms_hilo, dim3(M), dim3(256), st, (const float*)h->x, h->w.ln1 + (size_t)l * H, h->hl, hl_plane));
    if (gemm_launch(h->gemms[l * 4 + 0], st)) return CBX_ERR;
    if (h->cfg.kv_fp32 ? launch_attn<float>(h, l, st) : launch_attn<__nv_bfloat16>(h, l, st)) return CBX_ERR;
    if (gemm_launch(h->gemms[l * 4 + 1], st)) return CBX_ERR;
    CBX_CUDA(launch_pdl(k_rms_hilo, dim3(M), dim3(256), st, (const float*)h->x, h->w.ln2 + (size_t)l * H, h->hl, hl_plane));
    if (gemm_launch(h->gemms[l * 4 + 2], st)) return CBX_ERR;
    if (gemm_launch(h->gemms[l * 4 + 3], st)) return CBX_ERR;
    n += 7 + (h->S > 1 ? 1 : 0);
  }
  CBX_CUDA(launch_pdl(k_rms_hilo, dim3(M), dim3(256), st, (const float*)h->x, h->w.final_norm, h->hl, hl_plane));
  if (gemm_launch(h->gemms[L * 4], st)) return CBX_ERR;
  h->launches_per_step = n + 2;
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" int cbx_t3_batch_create(const CbxT3BatchConfig* cfg, const CbxT3Weights* w, cbx_t3b** out) {
  CBX_CHECK(cfg && w && out, "cbx_t3_batch_create: null argument");
  CBX_CHECK(w->wplain, "cbx_t3_batch_create: the batched engine needs the plain weight matrices (CbxT3Weights.wplain)");
  CBX_CHECK(cfg->n_layers >= 1 && cfg->n_layers <= 64, "cbx_t3_batch_create: n_layers %d out of range", cfg->n_layers);
  CBX_CHECK(cfg->n_slots >= 1 && cfg->n_slots <= 64, "cbx_t3_batch_create: n_slots %d out of range (1..64: 2 rows per slot, one 128-row tile)", cfg->n_slots);
  CBX_CHECK(cfg->max_pos >= 8 && cfg->max_new_cap >= 1, "cbx_t3_batch_create: max_pos / max_new_cap too small");
  CBX_ON_DEVICE(cfg->device);
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  CBX_CHECK(prop.major == 10, "cbx_t3_batch_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  cbx_t3b* h = new cbx_t3b();
  h->cfg = *cfg; h->w = *w; h->U = cfg->n_slots; h->M = 2 * cfg->n_slots;
  h->launches = 0; h->launches_per_step = 0; h->step_graph = nullptr; h->cap_stream = nullptr;
  const int M = h->M, U = h->U, L = cfg->n_layers;
  {   // KV splits: keep >= ~4 CTAs per SM in the attention grid at small batch
    const int ctas = M * NH;
    int S = ctas >= 512 ? 1 : (592 + ctas - 1) / ctas;
    h->S = S > MAX_SPLITS ? MAX_SPLITS : S;
    if (const char* v = getenv("CBX_T3B_SPLITS")) { int s = atoi(v); if (s >= 1 && s <= MAX_SPLITS) h->S = s; }
  }
  const size_t kvel = cfg->kv_fp32 ? 4 : 2;
  h->slot_stride = (long long)L * 2 * NH * cfg->max_pos * HD;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 1024); return o; };
  const size_t o_k = take((size_t)U * h->slot_stride * kvel), o_v = take((size_t)U * h->slot_stride * kvel);
  const size_t o_x = take((size_t)M * H * 4), o_qkv = take((size_t)M * 3072 * 4), o_log = take((size_t)M * VOCAB_LD * 4);
  const size_t o_part = take((size_t)M * NH * MAX_SPLITS * PART * 4);
  const size_t o_hl = take((size_t)2 * M * H * 2), o_act = take((size_t)2 * M * INTER * 2);
  const size_t o_st = take(sizeof(SlotState) * U), o_sp = take(sizeof(CbxSampleParams) * U), o_seen = take((size_t)U * SEEN_WORDS * 4);
  const size_t o_forced = take((size_t)U * cfg->max_new_cap * 4), o_tok = take((size_t)U * cfg->max_new_cap * 4);
  const size_t o_mask = take(U * 4), o_pfs = take(64);
  const size_t o_lout = cfg->keep_logits ? take((size_t)U * cfg->max_new_cap * 2 * VOCAB * 4) : 0;
  // prompt-prefill workspace: room for several utterances' prompts in one pass (cbx_t3_batch_prefill_many), at most 16 K rows
  {
    const long long one = 2LL * cfg->max_pos, all = one * U;
    h->pf_rows = all < 16384 ? all : (one > 16384 ? one : 16384);
  }
  const size_t o_ws = take(t3_prefill_workspace_bytes_rows((size_t)h->pf_rows));
  if (cudaMalloc(&h->slab, off) != cudaSuccess) {
    cbx_set_error("cbx_t3_batch_create: cudaMalloc of %zu bytes failed", off);
    delete h;
    return CBX_ERR;
  }
  cudaMemset(h->slab, 0, off);
  uint8_t* s = static_cast<uint8_t*>(h->slab);
  h->kc = s + o_k; h->vc = s + o_v;
  h->x = (float*)(s + o_x); h->qkv = (float*)(s + o_qkv); h->logits = (float*)(s + o_log); h->part = (float*)(s + o_part);
  h->hl = (__nv_bfloat16*)(s + o_hl); h->act = (__nv_bfloat16*)(s + o_act);
  h->st = (SlotState*)(s + o_st); h->sp = (CbxSampleParams*)(s + o_sp); h->seen = (unsigned*)(s + o_seen);
  h->forced = (int*)(s + o_forced); h->tokens = (int*)(s + o_tok); h->step_mask = (int*)(s + o_mask); h->pf_state = (int*)(s + o_pfs);
  h->logits_out = cfg->keep_logits ? (float*)(s + o_lout) : nullptr;
  h->prefill_ws = s + o_ws;
  h->host_st.resize(U);
  // every slot starts free: done = 1 (SlotState zero-filled means active = 0 already)
  // GEMM descriptors (tensor maps) of the step
  const TileCfg tc = tile_cfg();
  const __nv_bfloat16* wp = static_cast<const __nv_bfloat16*>(w->wplain);
  const size_t per_layer = (size_t)3072 * H + (size_t)H * H + (size_t)8192 * H + (size_t)H * INTER;
  const long long hl_plane = (long long)M * H, act_plane = (long long)M * INTER;
  h->gemms.resize(L * 4 + 1);
  int rc = CBX_OK;
  for (int l = 0; l < L && !rc; ++l) {
    const __nv_bfloat16* wqkv = wp + l * per_layer;
    const __nv_bfloat16* wo = wqkv + (size_t)3072 * H;
    const __nv_bfloat16* wgu = wo + (size_t)H * H;
    const __nv_bfloat16* wdn = wgu + (size_t)8192 * H;
    GemmEpi e{};
    e.out_f32 = h->qkv; e.ld_f32 = 3072;
    rc |= gemm_make_hilo_ex(&h->gemms[l * 4 + 0], h->hl, H, hl_plane, wqkv, M, 3072, 3072, H, e, tc.qkv_t, tc.qkv_s);
    e = GemmEpi{}; e.resid = h->x; e.ld_resid = H; e.out_f32 = h->x; e.ld_f32 = H;
    rc |= gemm_make_hilo_ex(&h->gemms[l * 4 + 1], h->hl, H, hl_plane, wo, M, H, H, H, e, tc.o_t, tc.o_s);
    e = GemmEpi{}; e.swiglu = 1; e.out_hilo = h->act; e.ld_hilo = INTER; e.hilo_plane = act_plane;
    rc |= gemm_make_hilo_ex(&h->gemms[l * 4 + 2], h->hl, H, hl_plane, wgu, M, 8192, 8192, H, e, 256, tc.gu_s);
    e = GemmEpi{}; e.resid = h->x; e.ld_resid = H; e.out_f32 = h->x; e.ld_f32 = H;
    rc |= gemm_make_hilo_ex(&h->gemms[l * 4 + 3], h->act, INTER, act_plane, wdn, M, H, H, INTER, e, tc.dn_t, tc.dn_s);
  }
  if (!rc) {
    GemmEpi e{};
    e.out_f32 = h->logits; e.ld_f32 = VOCAB_LD;
    rc |= gemm_make_hilo_ex(&h->gemms[L * 4], h->hl, H, hl_plane, wp + L * per_layer, M, VOCAB_LD, VOCAB, H, e, tc.hd_t, tc.hd_s);
  }
  if (rc) { cudaFree(h->slab); delete h; return CBX_ERR; }
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_t3_batch_destroy(cbx_t3b* h) {
  if (!h) return CBX_OK;
  if (h->step_graph) cudaGraphExecDestroy(h->step_graph);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

extern "C" int cbx_t3_batch_prefill(cbx_t3b* h, int slot, const float* embeds, int P, const CbxSampleParams* sp,
                                    int max_new_tokens, const int32_t* forced_tokens, void* stream) {
  CBX_CHECK(h && embeds && sp, "cbx_t3_batch_prefill: null argument");
  CBX_CHECK(slot >= 0 && slot < h->U, "cbx_t3_batch_prefill: slot %d out of range (%d slots)", slot, h->U);
  CBX_CHECK(max_new_tokens >= 1 && max_new_tokens <= h->cfg.max_new_cap, "cbx_t3_batch_prefill: max_new_tokens %d exceeds the slot's token capacity %d",
            max_new_tokens, h->cfg.max_new_cap);
  CBX_CHECK(P >= 1 && P + max_new_tokens <= h->cfg.max_pos, "cbx_t3_batch_prefill: %d prompt + %d new positions exceed the slot's KV slab (max_pos %d)",
            P, max_new_tokens, h->cfg.max_pos);
  CBX_CHECK(h->cfg.speech_pos_rows <= 0 || max_new_tokens + 1 <= h->cfg.speech_pos_rows,
            "cbx_t3_batch_prefill: max_new_tokens exceeds the learned speech position table");
  CBX_CHECK(sp->temperature > 0.f && sp->repetition_penalty > 0.f, "cbx_t3_batch_prefill: temperature and repetition_penalty must be > 0");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->cfg.device);
  const size_t kvel = h->cfg.kv_fp32 ? 4 : 2;
  T3PrefillArgs a{};
  a.embeds = embeds; a.B = 2; a.P = P; a.n_layers = h->cfg.n_layers; a.max_pos = h->cfg.max_pos; a.kv_fp32 = h->cfg.kv_fp32;
  a.wplain = h->w.wplain; a.ln1 = h->w.ln1; a.ln2 = h->w.ln2; a.final_norm = h->w.final_norm;
  a.rope_cos = h->w.rope_cos; a.rope_sin = h->w.rope_sin;
  a.kcache = static_cast<uint8_t*>(h->kc) + (size_t)slot * h->slot_stride * kvel;
  a.vcache = static_cast<uint8_t*>(h->vc) + (size_t)slot * h->slot_stride * kvel;
  a.logits = h->logits + (size_t)(2 * slot) * VOCAB_LD; a.logits_ld = VOCAB_LD;
  a.state = h->pf_state; a.workspace = h->prefill_ws; a.ws_rows = h->pf_rows; a.al_layer = -1;
  if (t3_prefill_gemm(a, st)) return CBX_ERR;
  CBX_CUDA(cudaMemcpyAsync(h->sp + slot, sp, sizeof(CbxSampleParams), cudaMemcpyHostToDevice, st));
  if (forced_tokens)
    CBX_CUDA(cudaMemcpyAsync(h->forced + (size_t)slot * h->cfg.max_new_cap, forced_tokens, (size_t)max_new_tokens * 4,
                             cudaMemcpyDeviceToDevice, st));
  k_b_slot_init<<<1, 256, 0, st>>>(h->st + slot, h->seen + (size_t)slot * SEEN_WORDS, P, max_new_tokens, forced_tokens ? 1 : 0);
  CBX_CUDA(cudaGetLastError());
  h->launches += 1 + h->cfg.n_layers * 8 + 3;
  return CBX_OK;
}

extern "C" int cbx_t3_batch_prefill_many(cbx_t3b* h, int slot0, int n, const float* embeds, int P, const CbxSampleParams* sp,
                                         int max_new_tokens, const int32_t* forced_tokens, void* stream) {
  CBX_CHECK(h && embeds && sp, "cbx_t3_batch_prefill_many: null argument");
  CBX_CHECK(n >= 1 && slot0 >= 0 && slot0 + n <= h->U, "cbx_t3_batch_prefill_many: slots [%d, %d) out of range (%d slots)", slot0, slot0 + n, h->U);
  CBX_CHECK(max_new_tokens >= 1 && max_new_tokens <= h->cfg.max_new_cap, "cbx_t3_batch_prefill_many: max_new_tokens %d exceeds the slot's token capacity %d",
            max_new_tokens, h->cfg.max_new_cap);
  CBX_CHECK(P >= 1 && P + max_new_tokens <= h->cfg.max_pos, "cbx_t3_batch_prefill_many: %d prompt + %d new positions exceed the slot's KV slab (max_pos %d)",
            P, max_new_tokens, h->cfg.max_pos);
  CBX_CHECK(h->cfg.speech_pos_rows <= 0 || max_new_tokens + 1 <= h->cfg.speech_pos_rows,
            "cbx_t3_batch_prefill_many: max_new_tokens exceeds the learned speech position table");
  for (int u = 0; u < n; ++u)
    CBX_CHECK(sp[u].temperature > 0.f && sp[u].repetition_penalty > 0.f, "cbx_t3_batch_prefill_many: temperature and repetition_penalty must be > 0 (request %d)", u);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->cfg.device);
  const size_t kvel = h->cfg.kv_fp32 ? 4 : 2;
  const int per_pass = (int)(h->pf_rows / (2LL * P));        // utterances whose prompts fit the workspace at once (>= 1: P <= max_pos)
  for (int u0 = 0; u0 < n; u0 += per_pass) {
    const int nb = n - u0 < per_pass ? n - u0 : per_pass, slot = slot0 + u0;
    T3PrefillArgs a{};
    a.embeds = embeds + (size_t)u0 * 2 * P * H; a.B = 2 * nb; a.P = P; a.n_layers = h->cfg.n_layers; a.max_pos = h->cfg.max_pos; a.kv_fp32 = h->cfg.kv_fp32;
    a.wplain = h->w.wplain; a.ln1 = h->w.ln1; a.ln2 = h->w.ln2; a.final_norm = h->w.final_norm;
    a.rope_cos = h->w.rope_cos; a.rope_sin = h->w.rope_sin;
    a.kcache = static_cast<uint8_t*>(h->kc) + (size_t)slot * h->slot_stride * kvel;
    a.vcache = static_cast<uint8_t*>(h->vc) + (size_t)slot * h->slot_stride * kvel;
    a.logits = h->logits + (size_t)(2 * slot) * VOCAB_LD; a.logits_ld = VOCAB_LD;
    a.state = h->pf_state; a.workspace = h->prefill_ws; a.ws_rows = h->pf_rows; a.al_layer = -1;
    a.nseq = nb; a.kv_seq_stride = h->slot_stride;
    if (t3_prefill_gemm(a, st)) return CBX_ERR;
    h->launches += 1 + h->cfg.n_layers * 8 + 2;
  }
  CBX_CUDA(cudaMemcpyAsync(h->sp + slot0, sp, sizeof(CbxSampleParams) * n, cudaMemcpyHostToDevice, st));
  if (forced_tokens)
    CBX_CUDA(cudaMemcpy2DAsync(h->forced + (size_t)slot0 * h->cfg.max_new_cap, (size_t)h->cfg.max_new_cap * 4, forced_tokens,
                               (size_t)max_new_tokens * 4, (size_t)max_new_tokens * 4, n, cudaMemcpyDeviceToDevice, st));
  for (int u = 0; u < n; ++u)
    k_b_slot_init<<<1, 256, 0, st>>>(h->st + slot0 + u, h->seen + (size_t)(slot0 + u) * SEEN_WORDS, P, max_new_tokens, forced_tokens ? 1 : 0);
  CBX_CUDA(cudaGetLastError());
  h->launches += n;
  return CBX_OK;
}

extern "C" int cbx_t3_decode_batch(cbx_t3b* h, const int32_t* slots, int n_slots, int n_steps, void* stream) {
  CBX_CHECK(h, "cbx_t3_decode_batch: null handle");
  CBX_CHECK(n_steps >= 1, "cbx_t3_decode_batch: n_steps must be >= 1");
  CBX_CHECK(slots == nullptr || (n_slots >= 1 && n_slots <= h->U), "cbx_t3_decode_batch: n_slots %d out of range", n_slots);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->cfg.device);
  std::vector<int> mask(h->U, slots ? 0 : 1);
  if (slots)
    for (int i = 0; i < n_slots; ++i) {
      CBX_CHECK(slots[i] >= 0 && slots[i] < h->U, "cbx_t3_decode_batch: slot %d out of range (%d slots)", slots[i], h->U);
      mask[slots[i]] = 1;
    }
  CBX_CUDA(cudaMemcpyAsync(h->step_mask, mask.data(), h->U * sizeof(int), cudaMemcpyHostToDevice, st));   // pageable: staged before return
  static const bool use_graph = [] { const char* v = getenv("CBX_T3B_GRAPH"); return !(v && v[0] == '0'); }();
  if (use_graph && !h->step_graph) {
    if (!h->cap_stream) CBX_CUDA(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
    cudaGraph_t graph = nullptr;
    CBX_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
    const int rc = step_body(h, h->cap_stream);
    const cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
    if (rc || ce != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      cbx_set_error("cbx_t3_decode_batch: graph capture failed (%s)", cudaGetErrorString(ce));
      return CBX_ERR;
    }
    const cudaError_t ie = cudaGraphInstantiate(&h->step_graph, graph, 0);
    cudaGraphDestroy(graph);
    CBX_CHECK(ie == cudaSuccess, "cbx_t3_decode_batch: cudaGraphInstantiate failed (%s)", cudaGetErrorString(ie));
  }
  for (int i = 0; i < n_steps; ++i) {
    if (use_graph) CBX_CUDA(cudaGraphLaunch(h->step_graph, st));
    else if (step_body(h, st)) return CBX_ERR;
  }
  h->launches += (long long)n_steps * h->launches_per_step;
  return CBX_OK;
}

extern "C" int cbx_t3_batch_poll(cbx_t3b* h, int32_t* n_gen, int32_t* done, int32_t* pos, void* stream) {
  CBX_CHECK(h, "cbx_t3_batch_poll: null handle");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->cfg.device);
  CBX_CUDA(cudaMemcpyAsync(h->host_st.data(), h->st, sizeof(SlotState) * h->U, cudaMemcpyDeviceToHost, st));
  CBX_CUDA(cudaStreamSynchronize(st));
  for (int u = 0; u < h->U; ++u) {
    const SlotState& s = h->host_st[u];
    if (n_gen) n_gen[u] = s.n_gen;
    if (done) done[u] = s.active ? s.done : -1;      // -1: free slot
    if (pos) pos[u] = s.pos + (s.fwd ? 1 : 0);
  }
  return CBX_OK;
}

extern "C" int cbx_t3_batch_tokens(cbx_t3b* h, int slot, int32_t* tokens_out, int n, void* stream) {
  CBX_CHECK(h && tokens_out, "cbx_t3_batch_tokens: null argument");
  CBX_CHECK(slot >= 0 && slot < h->U && n >= 0 && n <= h->cfg.max_new_cap, "cbx_t3_batch_tokens: slot %d / count %d out of range", slot, n);
  CBX_ON_DEVICE(h->cfg.device);
  CBX_CUDA(cudaMemcpyAsync(tokens_out, h->tokens + (size_t)slot * h->cfg.max_new_cap, (size_t)n * 4, cudaMemcpyDeviceToDevice,
                           static_cast<cudaStream_t>(stream)));
  return CBX_OK;
}

extern "C" int cbx_t3_batch_logits(cbx_t3b* h, int slot, float* logits_out, int n, void* stream) {
  CBX_CHECK(h && logits_out, "cbx_t3_batch_logits: null argument");
  CBX_CHECK(h->logits_out, "cbx_t3_batch_logits: the engine was created without keep_logits");
  CBX_CHECK(slot >= 0 && slot < h->U && n >= 0 && n <= h->cfg.max_new_cap, "cbx_t3_batch_logits: slot %d / count %d out of range", slot, n);
  CBX_ON_DEVICE(h->cfg.device);
  CBX_CUDA(cudaMemcpyAsync(logits_out, h->logits_out + (size_t)slot * h->cfg.max_new_cap * 2 * VOCAB, (size_t)n * 2 * VOCAB * 4,
                           cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return CBX_OK;
}

extern "C" int cbx_t3_batch_release(cbx_t3b* h, int slot, void* stream) {
  CBX_CHECK(h, "cbx_t3_batch_release: null handle");
  CBX_CHECK(slot >= 0 && slot < h->U, "cbx_t3_batch_release: slot %d out of range", slot);
  CBX_ON_DEVICE(h->cfg.device);
  k_b_slot_release<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(h->st + slot);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

extern "C" int cbx_t3_batch_info(cbx_t3b* h, int32_t* kv_splits, int64_t* launches, int64_t* launches_per_step) {
  CBX_CHECK(h, "cbx_t3_batch_info: null handle");
  if (kv_splits) *kv_splits = h->S;
  if (launches) *launches = h->launches;
  if (launches_per_step) *launches_per_step = h->launches_per_step;
  return CBX_OK;
}
