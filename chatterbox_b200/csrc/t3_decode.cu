// T3 (0.5B Llama) prefill + decode loop as ONE persistent cooperative kernel for sm_100a.
//
// Replaces the reference's per-token Python loop (src/chatterbox/models/t3/t3.py:319-367 and
// T3HuggingfaceBackend.forward, t3/inference/t3_hf_backend.py:73-116): ~600 library launches and a
// host sync per token become zero launches per token.  The step is HBM-bound (1.02 GB of bf16
// weights per token at batch 2), so the design is a weight-streaming pipeline:
//   * one CTA per SM (cooperative launch), 8 consumer warps + 1 producer warp;
//   * the producer walks this CTA's pre-packed, contiguous weight stream with 16 KB TMA bulk
//     copies (cp.async.bulk -> UBLKCP) into an 11-deep shared-memory ring guarded by mbarriers.
//     Weights do not depend on activations, so the producer never waits on a grid barrier: HBM
//     keeps streaming while the consumers synchronise between dependent phases;
//   * consumers multiply each 8-row x 1024-k unit against the (<= 8) activation rows with
//     mma.sync m16n8k16 (weights are the B operand; activations are split x = hi + lo in bf16 and
//     occupy A rows 0-7 / 8-15, so products carry ~16 mantissa bits and accumulate in fp32);
//   * dependent phases (qkv -> attention -> o_proj -> gate/up -> down) are separated by a
//     counter-based grid barrier; residual adds, RMSNorm, RoPE, SwiGLU, the split-KV combine, CFG,
//     logit processing, sampling, EOS and the next-token embedding are all fused into the phase
//     prologues/epilogues, so nothing but weights and the KV slab touches HBM.
// The same kernel runs chunked prefill (4 positions x 2 CFG rows per weight pass).
#include "common.cuh"
#include "t3_plan.h"
#include "../../include/cbx.h"

#include <cooperative_groups.h>
#include <math.h>
#include <stdarg.h>

namespace {

constexpr int NSLOT = 11;                 // ring depth (16 KB units)
constexpr int NCONS_WARPS = 8;
constexpr int NCONS = NCONS_WARPS * 32;   // consumer threads
constexpr int NTHREADS = NCONS + 32;      // + producer warp
constexpr int MAXS = 16;                  // max KV splits per (row, head) pair
constexpr int PART_STRIDE = 68;           // 64 o + m + l (+2 pad: rows stay 16 B aligned)
constexpr int STOP_SPEECH = 6562;
constexpr int VOCAB_PAD = 8200;
constexpr float RMS_EPS = 1e-5f;

constexpr int SMEM_RING = NSLOT * T3_UNIT_BYTES;
constexpr int SMEM_BARS = SMEM_RING;            // 2*NSLOT mbarriers
constexpr int SMEM_CTL = SMEM_BARS + 256;       // control words
constexpr int SMEM_SCRATCH = SMEM_CTL + 256;    // phase scratch (union)
constexpr int SCRATCH_BYTES = VOCAB_PAD * 4 + 1024;
constexpr int SMEM_TOTAL = SMEM_SCRATCH + SCRATCH_BYTES;

struct T3State {   // lives in global memory, persists between launches of one utterance
  int pos;         // KV cache length
  int n_gen;       // speech tokens generated so far (next learned speech position = n_gen + 1)
  int stop;        // EOS seen
  int err;         // device-side error (1 = watchdog timeout)
};

struct T3Params {
  const uint8_t* wstream; long long stride_bytes;
  const float *ln1, *ln2, *final_norm, *speech_emb, *speech_pos_emb, *rope_cos, *rope_sin;
  int n_layers, max_pos;
  void *kcache, *vcache;        // [layer][2][16][max_pos][64]
  float *qkv, *hres, *oproj, *act, *down, *logits, *part;
  unsigned* bar; T3State* st; unsigned char* seen;
  int mode;                     // 0 prefill, 1 decode
  int B;
  const float* embeds; int P;   // prefill
  CbxSampleParams sp; int max_new; const int* forced; int* tokens_out; int* n_out; float* logits_out;
  unsigned long long* prof; int prof_it;   // optional: CTA 0 stamps %globaltimer after every phase of iteration prof_it
};

struct Ctl {            // CTA-local control block in shared memory
  volatile int abort;   // set after a watchdog timeout: every later wait returns at once
  volatile int cons_done;
  volatile int cons_k;
  volatile int tok;
};

using namespace cbx;

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long WATCHDOG_NS = 4000000000ull;

__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory"); }

// Wait on an mbarrier phase with a watchdog; returns false if aborted.
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, Ctl* ctl, T3State* st) {
  if (mbar_try_wait(bar, parity)) return true;
  unsigned long long t0 = gtime_ns();
  unsigned spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (ctl->abort) return false;
    if ((++spins & 1023u) == 0 && gtime_ns() - t0 > WATCHDOG_NS) {
      ctl->abort = 1;
      atomicExch(&st->err, 1);
      return false;
    }
  }
  return true;
}

// Counter-based grid barrier (all CTAs co-resident: cooperative launch).  Consumer threads only.
__device__ __forceinline__ void grid_barrier(const T3Params& p, unsigned& target, Ctl* ctl) {
  target += gridDim.x;
  cbar();
  if (threadIdx.x == 0 && !ctl->abort) {
    __threadfence();
    red_release_add(p.bar, 1u);
    unsigned long long t0 = gtime_ns();
    unsigned spins = 0;
    while (ld_acquire_u(p.bar) < target) {
      if ((++spins & 255u) == 0) {
        if (gtime_ns() - t0 > WATCHDOG_NS) { ctl->abort = 1; atomicExch(&p.st->err, 1); break; }
        if (__ldcg(&p.st->err)) { ctl->abort = 1; break; }
      }
    }
  }
  cbar();
}

// ---- activation rows -> A fragments -------------------------------------------------------------
// Lane (g = lane>>2, tig = lane&3) owns activation row g = (pos pp = g>>1, CFG row bb = g&1) and, for
// each of the warp's four k32-blocks jj, the 8 consecutive k at warp*128 + jj*32 + tig*8.
struct Piece { const float* ptr; int sp, sb; };   // element strides over (pp, bb)

template <int NP>
__device__ __forceinline__ void load_x(float (&xs)[4][8], const Piece (&pc)[NP], int npieces, bool active, int pp,
                                       int bb, int kbase) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
#pragma unroll
    for (int e = 0; e < 8; ++e) xs[jj][e] = 0.f;
  if (!active) return;
#pragma unroll
  for (int q = 0; q < NP; ++q) {
    if (q < npieces) {
      const float* base = pc[q].ptr + (size_t)pp * pc[q].sp + (size_t)bb * pc[q].sb + kbase;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        float4 a = ldcg_f4(base + jj * 32), b = ldcg_f4(base + jj * 32 + 4);
        xs[jj][0] += a.x; xs[jj][1] += a.y; xs[jj][2] += a.z; xs[jj][3] += a.w;
        xs[jj][4] += b.x; xs[jj][5] += b.y; xs[jj][6] += b.z; xs[jj][7] += b.w;
      }
    }
  }
}

__device__ __forceinline__ void make_frags(uint32_t (&A)[4][8], const float (&xs)[4][8]) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    split_hi_lo(xs[jj][0], xs[jj][1], A[jj][0], A[jj][1]);   // step 0: a0 (hi), a1 (lo)  k = 2tig,2tig+1
    split_hi_lo(xs[jj][2], xs[jj][3], A[jj][2], A[jj][3]);   //         a2 (hi), a3 (lo)  k = 2tig+8,+9
    split_hi_lo(xs[jj][4], xs[jj][5], A[jj][4], A[jj][5]);   // step 1
    split_hi_lo(xs[jj][6], xs[jj][7], A[jj][6], A[jj][7]);
  }
}

// RMSNorm over the full 1024-wide row (reference: HF LlamaRMSNorm, eps 1e-5): the row is spread
// over the 4 tig lanes of 8 warps; partial sums of squares meet in shared memory.
__device__ __forceinline__ void rmsnorm_rows(float (&xs)[4][8], const float* __restrict__ lnw, int kbase, float* ssq,
                                             int warp, int lane) {
  float s = 0.f;
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
#pragma unroll
    for (int e = 0; e < 8; ++e) s += xs[jj][e] * xs[jj][e];
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  const int g = lane >> 2;
  if ((lane & 3) == 0) ssq[warp * 8 + g] = s;
  cbar();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) tot += ssq[w * 8 + g];
  const float scale = 1.0f / sqrtf(tot * (1.0f / T3_H) + RMS_EPS);
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    float4 a = __ldg(reinterpret_cast<const float4*>(lnw + kbase + jj * 32));
    float4 b = __ldg(reinterpret_cast<const float4*>(lnw + kbase + jj * 32 + 4));
    xs[jj][0] = a.x * (xs[jj][0] * scale); xs[jj][1] = a.y * (xs[jj][1] * scale);
    xs[jj][2] = a.z * (xs[jj][2] * scale); xs[jj][3] = a.w * (xs[jj][3] * scale);
    xs[jj][4] = b.x * (xs[jj][4] * scale); xs[jj][5] = b.y * (xs[jj][5] * scale);
    xs[jj][6] = b.z * (xs[jj][6] * scale); xs[jj][7] = b.w * (xs[jj][7] * scale);
  }
}

// ---- consume n weight units from the ring: red[j][warp][64] = this warp's k-slice partial of unit j
struct RingPos { int slot; uint32_t parity; int k; };

__device__ __forceinline__ void consume_units(int n, RingPos& rp, const uint32_t (&A)[4][8], float* red,
                                              const uint8_t* ring, uint64_t* full, uint64_t* empty, int warp, int lane,
                                              Ctl* ctl, T3State* st) {
  for (int j = 0; j < n; ++j) {
    mbar_wait(&full[rp.slot], rp.parity, ctl, st);
    const uint4* wp = reinterpret_cast<const uint4*>(ring + rp.slot * T3_UNIT_BYTES) + warp * 128 + lane;
    float c4[4] = {0.f, 0.f, 0.f, 0.f};
    uint4 w[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) w[jj] = wp[jj * 32];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      mma_bf16_16816(c4, A[jj][0], A[jj][1], A[jj][2], A[jj][3], w[jj].x, w[jj].y);
      mma_bf16_16816(c4, A[jj][4], A[jj][5], A[jj][6], A[jj][7], w[jj].z, w[jj].w);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[rp.slot]);
    reinterpret_cast<float2*>(red + (j * NCONS_WARPS + warp) * 64)[lane] = make_float2(c4[0] + c4[2], c4[1] + c4[3]);
    if (++rp.slot == NSLOT) { rp.slot = 0; rp.parity ^= 1u; }
    ++rp.k;
  }
}

__device__ __forceinline__ float red_sum(const float* red, int j, int o) {
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) s += red[(j * NCONS_WARPS + w) * 64 + o];
  return s;
}

}  // namespace

namespace {

// ---- KV slab element access ---------------------------------------------------------------------
template <typename KVT> struct KV;
template <> struct KV<__nv_bfloat16> {
  static __device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
    uint4 v = __ldcg(reinterpret_cast<const uint4*>(p));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  static __device__ __forceinline__ float round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }
};
template <> struct KV<float> {
  static __device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
    float4 a = __ldcg(reinterpret_cast<const float4*>(p)), b = __ldcg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
  static __device__ __forceinline__ float round(float x) { return x; }
  static __device__ __forceinline__ void store(float* p, float x) { *p = x; }
};

// online-softmax state of one query over the keys a lane group has seen
struct Osm { float m, l, o[8]; };
__device__ __forceinline__ void osm_init(Osm& s) {
  s.m = -INFINITY; s.l = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) s.o[e] = 0.f;
}
__device__ __forceinline__ void osm_update(Osm& s, float score, const float (&v)[8]) {
  float mn = fmaxf(s.m, score);
  float corr = __expf(s.m - mn);     // exp(-inf) = 0 on first use
  float pe = __expf(score - mn);
  s.l = s.l * corr + pe;
#pragma unroll
  for (int e = 0; e < 8; ++e) s.o[e] = s.o[e] * corr + pe * v[e];
  s.m = mn;
}
__device__ __forceinline__ void osm_merge_xor(Osm& s, int mask) {
  float m2 = __shfl_xor_sync(0xffffffffu, s.m, mask);
  float l2 = __shfl_xor_sync(0xffffffffu, s.l, mask);
  float mn = fmaxf(s.m, m2);
  float c1 = (s.m == -INFINITY) ? 0.f : __expf(s.m - mn);
  float c2 = (m2 == -INFINITY) ? 0.f : __expf(m2 - mn);
  s.l = s.l * c1 + l2 * c2;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    float o2 = __shfl_xor_sync(0xffffffffu, s.o[e], mask);
    s.o[e] = s.o[e] * c1 + o2 * c2;
  }
  s.m = mn;
}

// ---- phase B: RoPE + KV append + split-KV attention over the slab ------------------------------
// (reference: HF LlamaAttention eager path: softmax(q k^T / 8 + causal) v in fp32; rotate_half RoPE)
// CTA c < S*npairs handles pair = c % npairs (CFG row b, head) and KV split c / npairs.
// Output: part[pp][pair][split][0..63] = sum_i exp(s_i - m) v_i, [64] = m, [65] = l.
template <typename KVT>
__device__ void attn_phase(const T3Params& p, int layer, int pos0, int npos, float* scr, int warp, int lane) {
  const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x;
  const int npairs = p.B * T3_NH;
  int S = G / npairs; if (S > MAXS) S = MAXS;
  if (c >= S * npairs) return;
  const int pair = c % npairs, split = c / npairs, b = pair / T3_NH, head = pair % T3_NH;
  float* q_s = scr;                       // [4][64] (pre-scaled by 1/8)
  float* k_new = scr + 256;               // [4][64]
  float* v_new = scr + 512;               // [4][64]
  float* wpart = scr + 768;               // [8 warps][4][66]
  KVT* kc = reinterpret_cast<KVT*>(p.kcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;
  KVT* vc = reinterpret_cast<KVT*>(p.vcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;

  if (tid < npos * 64) {
    const int pp = tid >> 6, d = tid & 63;
    const float* row = p.qkv + (size_t)(pp * 2 + b) * 3072 + head * 64;
    const float cs = __ldg(p.rope_cos + (size_t)(pos0 + pp) * 32 + (d & 31));
    const float sn = __ldg(p.rope_sin + (size_t)(pos0 + pp) * 32 + (d & 31));
    const float q = ldcg_f(row + d), qo = ldcg_f(row + (d ^ 32));
    const float k = ldcg_f(row + 1024 + d), ko = ldcg_f(row + 1024 + (d ^ 32));
    const float v = ldcg_f(row + 2048 + d);
    const float qr = q * cs + (d < 32 ? -qo : qo) * sn;
    const float kr = KV<KVT>::round(k * cs + (d < 32 ? -ko : ko) * sn);
    const float vr = KV<KVT>::round(v);
    q_s[pp * 64 + d] = qr * 0.125f;
    k_new[pp * 64 + d] = kr;
    v_new[pp * 64 + d] = vr;
    if (split == S - 1) {
      KV<KVT>::store(kc + (size_t)(pos0 + pp) * 64 + d, kr);
      KV<KVT>::store(vc + (size_t)(pos0 + pp) * 64 + d, vr);
    }
  }
  cbar();

  const int lg = lane >> 3, sub = lane & 7;
  float qv[T3_MAX_NPOS][8];
  Osm st[T3_MAX_NPOS];
#pragma unroll
  for (int pp = 0; pp < T3_MAX_NPOS; ++pp) {
    osm_init(st[pp]);
#pragma unroll
    for (int e = 0; e < 8; ++e) qv[pp][e] = (pp < npos) ? q_s[pp * 64 + sub * 8 + e] : 0.f;
  }
  const int chunk = (pos0 + S - 1) / S;
  const int lo = split * chunk;
  const int hi = min(pos0, lo + chunk);
  for (int base = lo + warp * 4; base < hi; base += NCONS_WARPS * 4) {
    const int i = base + lg;
    const bool valid = i < hi;
    float kf[8], vf[8];
    if (valid) {
      KV<KVT>::load8(kc + (size_t)i * 64 + sub * 8, kf);
      KV<KVT>::load8(vc + (size_t)i * 64 + sub * 8, vf);
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) { kf[e] = 0.f; vf[e] = 0.f; }
    }
#pragma unroll
    for (int pp = 0; pp < T3_MAX_NPOS; ++pp) {
      if (pp < npos) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s += qv[pp][e] * kf[e];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        if (valid) osm_update(st[pp], s, vf);
      }
    }
  }
  // keys of this chunk of positions itself (causal inside the chunk): last split, warp 0, lane group = key
  if (split == S - 1 && warp == 0) {
    const bool valid = lg < npos;
    float kf[8], vf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      kf[e] = valid ? k_new[lg * 64 + sub * 8 + e] : 0.f;
      vf[e] = valid ? v_new[lg * 64 + sub * 8 + e] : 0.f;
    }
#pragma unroll
    for (int pp = 0; pp < T3_MAX_NPOS; ++pp) {
      if (pp < npos) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s += qv[pp][e] * kf[e];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        if (valid && lg <= pp) osm_update(st[pp], s, vf);
      }
    }
  }
  // merge the 4 lane groups, then the 8 warps
#pragma unroll
  for (int pp = 0; pp < T3_MAX_NPOS; ++pp) {
    if (pp < npos) {
      osm_merge_xor(st[pp], 8);
      osm_merge_xor(st[pp], 16);
      if (lg == 0) {
        float* w = wpart + (warp * T3_MAX_NPOS + pp) * PART_STRIDE;
#pragma unroll
        for (int e = 0; e < 8; ++e) w[sub * 8 + e] = st[pp].o[e];
        if (sub == 0) { w[64] = st[pp].m; w[65] = st[pp].l; }
      }
    }
  }
  cbar();
  if (tid < npos * 64) {
    const int pp = tid >> 6, d = tid & 63;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NCONS_WARPS; ++w) M = fmaxf(M, wpart[(w * T3_MAX_NPOS + pp) * PART_STRIDE + 64]);
    float L = 0.f, O = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < NCONS_WARPS; ++w) {
        const float* wp = wpart + (w * T3_MAX_NPOS + pp) * PART_STRIDE;
        const float cw = (wp[64] == -INFINITY) ? 0.f : __expf(wp[64] - M);
        L += cw * wp[65];
        O += cw * wp[d];
      }
    }
    float* out = p.part + ((size_t)(pp * 32 + pair) * MAXS + split) * PART_STRIDE;
    out[d] = O;
    if (d == 0) { out[64] = M; out[65] = L; }
  }
}

// ---- phase C prologue: combine the KV splits into the attention output rows --------------------
__device__ __forceinline__ void load_attn_x(float (&xs)[4][8], const T3Params& p, bool active, int pp, int bb, int warp,
                                            int tig) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
#pragma unroll
    for (int e = 0; e < 8; ++e) xs[jj][e] = 0.f;
  if (!active) return;
  const int npairs = p.B * T3_NH;
  int S = gridDim.x / npairs; if (S > MAXS) S = MAXS;
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    const int head = 2 * warp + (jj >> 1), d0 = (jj & 1) * 32 + tig * 8;
    const float* base = p.part + ((size_t)(pp * 32 + bb * T3_NH + head) * MAXS) * PART_STRIDE;
    float M = -INFINITY;
    for (int s = 0; s < S; ++s) M = fmaxf(M, ldcg_f(base + s * PART_STRIDE + 64));
    float L = 0.f;
    for (int s = 0; s < S; ++s) {
      const float ms = ldcg_f(base + s * PART_STRIDE + 64);
      const float cw = (ms == -INFINITY) ? 0.f : __expf(ms - M);
      L += cw * ldcg_f(base + s * PART_STRIDE + 65);
      const float4 a = ldcg_f4(base + s * PART_STRIDE + d0), bq = ldcg_f4(base + s * PART_STRIDE + d0 + 4);
      xs[jj][0] += cw * a.x; xs[jj][1] += cw * a.y; xs[jj][2] += cw * a.z; xs[jj][3] += cw * a.w;
      xs[jj][4] += cw * bq.x; xs[jj][5] += cw * bq.y; xs[jj][6] += cw * bq.z; xs[jj][7] += cw * bq.w;
    }
    const float inv = 1.0f / L;
#pragma unroll
    for (int e = 0; e < 8; ++e) xs[jj][e] *= inv;
  }
}

}  // namespace

namespace {

// ---- sampling phase (CTA 0): t3.py:320-356 ------------------------------------------------------
__device__ __forceinline__ float block_sum(float v, float* rs, int warp, int lane) {
  v = warp_sum(v);
  cbar();
  if (lane == 0) rs[warp] = v;
  cbar();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) t += rs[w];
  return t;
}

// Philox4x32-10 -> one uniform in [0,1)
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

// order-preserving float <-> uint map for bisection on values
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Processes the pending logits of the CFG pair and emits token `it`.  All NCONS threads of CTA 0.
__device__ void sample_phase(const T3Params& p, int it, int n_gen, float* sl, Ctl* ctl, int warp, int lane) {
  const int tid = threadIdx.x;
  float* rs = sl + VOCAB_PAD;          // reduction scratch (64 floats)
  int* ri = reinterpret_cast<int*>(rs + 32);
  const CbxSampleParams& sp = p.sp;
  const bool cfg = sp.cfg_weight > 0.f && p.B == 2;
  // 1. CFG, temperature, repetition penalty; running max (first index on ties)
  float lmax = -INFINITY; int imax = 0x7fffffff;
  for (int i = tid; i < T3_VOCAB; i += NCONS) {
    const float c = ldcg_f(p.logits + i);
    const float u = (p.B == 2) ? ldcg_f(p.logits + T3_VOCAB + i) : 0.f;
    if (p.logits_out) {
      p.logits_out[((size_t)it * p.B) * T3_VOCAB + i] = c;
      if (p.B == 2) p.logits_out[((size_t)it * p.B + 1) * T3_VOCAB + i] = u;
    }
    float x = cfg ? c + sp.cfg_weight * (c - u) : c;
    if (sp.temperature != 1.0f) x = x / sp.temperature;
    if (__ldcg(p.seen + i)) x = (x < 0.f) ? x * sp.repetition_penalty : x / sp.repetition_penalty;
    sl[i] = x;
    if (x > lmax) { lmax = x; imax = i; }
  }
  // block arg-max
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float m2 = __shfl_xor_sync(0xffffffffu, lmax, o);
    int i2 = __shfl_xor_sync(0xffffffffu, imax, o);
    if (m2 > lmax || (m2 == lmax && i2 < imax)) { lmax = m2; imax = i2; }
  }
  cbar();
  if (lane == 0) { rs[warp] = lmax; ri[warp] = imax; }
  cbar();
  float M = rs[0]; int iM = ri[0];
#pragma unroll
  for (int w = 1; w < NCONS_WARPS; ++w)
    if (rs[w] > M || (rs[w] == M && ri[w] < iM)) { M = rs[w]; iM = ri[w]; }
  int tok = iM;
  if (!sp.greedy && p.forced == nullptr) {
    // 2. min-p (HF MinPLogitsWarper): drop p_i < min_p * p_max  <=>  exp(x_i - M) < min_p; top-1 kept
    float zloc = 0.f;
    for (int i = tid; i < T3_VOCAB; i += NCONS) {
      float e = __expf(sl[i] - M);
      if (sp.min_p > 0.f && e < sp.min_p && i != iM) { sl[i] = -INFINITY; e = 0.f; }
      zloc += e;
    }
    float Z = block_sum(zloc, rs, warp, lane);
    // 3. top-p (HF TopPLogitsWarper): drop tokens whose ascending cumulative prob <= 1 - top_p
    if (sp.top_p < 1.0f) {
      const float thr = (1.0f - sp.top_p) * Z;
      unsigned lo = 0u, hi = f2ord(M);       // largest ordinal t with S(t) = sum_{x_i <= t} e_i <= thr
      for (int itb = 0; itb < 32 && lo < hi; ++itb) {
        const unsigned mid = lo + (hi - lo + 1) / 2;
        float s = 0.f;
        for (int i = tid; i < T3_VOCAB; i += NCONS) {
          const float x = sl[i];
          if (x != -INFINITY && f2ord(x) <= mid) s += __expf(x - M);
        }
        s = block_sum(s, rs, warp, lane);
        if (s <= thr) lo = mid; else hi = mid - 1;
      }
      zloc = 0.f;
      for (int i = tid; i < T3_VOCAB; i += NCONS) {
        const float x = sl[i];
        if (x != -INFINITY && f2ord(x) <= lo && i != iM) sl[i] = -INFINITY;
        else if (x != -INFINITY) zloc += __expf(x - M);
      }
      Z = block_sum(zloc, rs, warp, lane);
    }
    // 4. multinomial: inverse CDF in index order (contiguous 33-wide chunk per thread)
    const float target = philox_uniform(sp.seed, (unsigned)(n_gen)) * Z;
    constexpr int CH = (T3_VOCAB + NCONS - 1) / NCONS;
    const int i0 = tid * CH, i1 = min(T3_VOCAB, i0 + CH);
    float loc = 0.f;
    for (int i = i0; i < i1; ++i) loc += (sl[i] == -INFINITY) ? 0.f : __expf(sl[i] - M);
    // inclusive scan of per-thread sums across the block
    float inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { float t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    cbar();
    if (lane == 31) rs[warp] = inc;
    if (tid == 0) ri[16] = iM;      // fallback when rounding leaves target beyond the last bin
    cbar();
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
    cbar();
    tok = ri[16];
  }
  if (p.forced) tok = p.forced[it];
  if (tid == 0) {
    p.tokens_out[it] = tok;
    p.seen[tok] = 1;
    if (sp.stop_on_eos && tok == STOP_SPEECH) { p.st->stop = 1; }
    ctl->tok = tok;
  }
  // 5. next-token embedding: speech_emb[tok] + speech_pos_emb[n_gen + 1]  (t3.py:351-356), both CFG rows
  {
    const float4 e = __ldg(reinterpret_cast<const float4*>(p.speech_emb + (size_t)tok * T3_H) + tid);
    const float4 q = __ldg(reinterpret_cast<const float4*>(p.speech_pos_emb + (size_t)(n_gen + 1) * T3_H) + tid);
    const float4 h = make_float4(e.x + q.x, e.y + q.y, e.z + q.z, e.w + q.w);
    float* dst = p.hres + (size_t)1 * (T3_MAX_NPOS * 2 * T3_H);   // hres[1][pp=0][b]
    for (int b = 0; b < p.B; ++b) reinterpret_cast<float4*>(dst + b * T3_H)[tid] = h;
  }
}

}  // namespace

namespace {

// ---- the megakernel -----------------------------------------------------------------------------
template <typename KVT>
__global__ void __launch_bounds__(NTHREADS, 1) t3_megakernel(const T3Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SMEM_BARS);
  uint64_t* empty = full + NSLOT;
  Ctl* ctl = reinterpret_cast<Ctl*>(smem + SMEM_CTL);
  float* scr = reinterpret_cast<float*>(smem + SMEM_SCRATCH);
  float* red = scr;                       // [MAXU][8][64]
  float* ssq = scr + T3_MAXU * 8 * 64;    // [8][8]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, c = blockIdx.x;
  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NCONS_WARPS); }
    ctl->abort = 0; ctl->cons_done = 0; ctl->cons_k = 0; ctl->tok = 0;
    mbar_fence_init();
  }
  __syncthreads();

  const int L = p.n_layers;
  const int upl = t3_plan_units_per_layer(c, G);
  const int nhead = t3_plan_count(PH_HEAD, c, G);
  const int n_iters = (p.mode == 0) ? (p.P + T3_MAX_NPOS - 1) / T3_MAX_NPOS : p.max_new;

  if (warp == NCONS_WARPS) {
    // ===== producer warp: stream this CTA's weight units through the ring =====
    const uint64_t policy = l2_policy_evict_first();
    const uint8_t* base = p.wstream + (size_t)c * p.stride_bytes;
    int slot = 0; uint32_t parity = 1; int k = 0;
    bool stopped = false;
    for (int it = 0; it < n_iters && !stopped; ++it) {
      const bool head = (p.mode == 1) || (it == n_iters - 1);
      const int n_units = L * upl + (head ? nhead : 0);
      const int head_skip = head ? 0 : 0;
      (void)head_skip;
      for (int u = 0; u < n_units; ++u) {
        // wait for the slot to be free; give up if the consumers have finished (EOS / abort)
        unsigned long long t0 = 0; unsigned spins = 0;
        while (!mbar_try_wait(&empty[slot], parity)) {
          if (ctl->cons_done || ctl->abort) { stopped = true; break; }
          if ((++spins & 1023u) == 0) {
            if (t0 == 0) t0 = gtime_ns();
            else if (gtime_ns() - t0 > WATCHDOG_NS) { ctl->abort = 1; atomicExch(&p.st->err, 1); stopped = true; break; }
          }
        }
        if (stopped) break;
        if (lane == 0) {
          mbar_arrive_expect_tx(&full[slot], T3_UNIT_BYTES);
          tma_bulk_g2s(ring + slot * T3_UNIT_BYTES, base + (size_t)u * T3_UNIT_BYTES, T3_UNIT_BYTES, &full[slot], policy);
        }
        __syncwarp();
        if (++slot == NSLOT) { slot = 0; parity ^= 1u; }
        ++k;
      }
    }
    // drain: every issued copy must have landed before the CTA may exit
    {
      unsigned long long t0 = gtime_ns(); unsigned spins = 0;
      while (!ctl->cons_done) {
        if ((++spins & 1023u) == 0 && gtime_ns() - t0 > 4 * WATCHDOG_NS) break;
      }
      __threadfence_block();
      const int ck = ctl->cons_k;
      for (int kk = ck; kk < k; ++kk) mbar_wait(&full[kk % NSLOT], (kk / NSLOT) & 1, ctl, p.st);
    }
  } else {
    // ===== consumer warps =====
    const int g = lane >> 2, tig = lane & 3, pp = g >> 1, bb = g & 1;
    RingPos rp{0, 0u, 0};
    unsigned bar_target = 0;
    int pos0 = __ldcg(&p.st->pos);
    int n_gen = __ldcg(&p.st->n_gen);
    int n_done = 0;
    float xs[4][8];
    uint32_t A[4][8];
    const int HS = T3_MAX_NPOS * 2 * T3_H;   // one hres buffer
    int prof_k = 0;
#define T3_PROF() do { if (p.prof && c == 0 && tid == 0 && it == p.prof_it) p.prof[prof_k++] = gtime_ns(); } while (0)

    for (int it = 0; it < n_iters; ++it) {
      int npos = 1;
      if (p.mode == 1) {
        T3_PROF();
        if (c == 0) sample_phase(p, it, n_gen, scr, ctl, warp, lane);
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
        ++n_gen; ++n_done;
        if (__ldcg(&p.st->stop) || ctl->abort) break;
      } else {
        npos = min(T3_MAX_NPOS, p.P - it * T3_MAX_NPOS);
      }
      const bool active = (g < 2 * npos) && (bb < p.B);
      const int kw = warp * 128 + tig * 8;

      for (int l = 0; l < L; ++l) {
        // ---------- phase A: x_l = residual; RMSNorm; fused q|k|v GEMV ----------
        {
          Piece pc[6];
          int np;
          if (l == 0) {
            if (p.mode == 0) pc[0] = Piece{p.embeds + (size_t)it * T3_MAX_NPOS * T3_H, T3_H, p.P * T3_H};
            else pc[0] = Piece{p.hres + HS, 2 * T3_H, T3_H};
            np = 1;
          } else {
            pc[0] = Piece{p.hres + ((l - 1) & 1) * HS, 2 * T3_H, T3_H};
            pc[1] = Piece{p.oproj, 2 * T3_H, T3_H};
            for (int s = 0; s < 4; ++s) pc[2 + s] = Piece{p.down + s * HS, 2 * T3_H, T3_H};
            np = 6;
          }
          load_x<6>(xs, pc, np, active, pp, bb, kw);
          if (c == 0 && active) {
            float* dst = p.hres + (l & 1) * HS + (pp * 2 + bb) * T3_H + kw;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
              *reinterpret_cast<float4*>(dst + jj * 32) = make_float4(xs[jj][0], xs[jj][1], xs[jj][2], xs[jj][3]);
              *reinterpret_cast<float4*>(dst + jj * 32 + 4) = make_float4(xs[jj][4], xs[jj][5], xs[jj][6], xs[jj][7]);
            }
          }
          rmsnorm_rows(xs, p.ln1 + (size_t)l * T3_H, kw, ssq, warp, lane);
          make_frags(A, xs);
          const int n = t3_plan_count(PH_QKV, c, G);
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          cbar();
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (gg < 2 * npos && (gg & 1) < p.B)
              p.qkv[(size_t)gg * 3072 + t3_plan_unit(PH_QKV, c, G, j) * 8 + nn] = red_sum(red, j, o);
          }
        }
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
        // ---------- phase B: RoPE, KV append, split-KV attention ----------
        attn_phase<KVT>(p, l, pos0, npos, scr, warp, lane);
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
        // ---------- phase C: o_proj ----------
        {
          load_attn_x(xs, p, active, pp, bb, warp, tig);
          make_frags(A, xs);
          const int n = t3_plan_count(PH_O, c, G);
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          cbar();
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (gg < 2 * npos && (gg & 1) < p.B)
              p.oproj[(size_t)gg * T3_H + t3_plan_unit(PH_O, c, G, j) * 8 + nn] = red_sum(red, j, o);
          }
        }
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
        // ---------- phase D: post-attention RMSNorm; gate/up GEMV; SwiGLU ----------
        {
          Piece pc[2] = {Piece{p.hres + (l & 1) * HS, 2 * T3_H, T3_H}, Piece{p.oproj, 2 * T3_H, T3_H}};
          load_x<2>(xs, pc, 2, active, pp, bb, kw);
          rmsnorm_rows(xs, p.ln2 + (size_t)l * T3_H, kw, ssq, warp, lane);
          make_frags(A, xs);
          const int n = t3_plan_count(PH_GU, c, G);
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          cbar();
          for (int idx = tid; idx < n * 32; idx += NCONS) {
            const int j = idx >> 5, o2 = idx & 31, gg = o2 >> 2, nn = o2 & 3;
            if (gg < 2 * npos && (gg & 1) < p.B) {
              const float gate = red_sum(red, j, gg * 8 + nn), up = red_sum(red, j, gg * 8 + 4 + nn);
              const float silu = gate / (1.0f + __expf(-gate));
              p.act[(size_t)gg * T3_INTER + t3_plan_unit(PH_GU, c, G, j) * 4 + nn] = silu * up;
            }
          }
        }
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
        // ---------- phase E: down_proj (k-slice s of this CTA) ----------
        {
          const int n = t3_plan_count(PH_DOWN, c, G);
          if (n > 0) {
            const int u0 = t3_plan_unit(PH_DOWN, c, G, 0), s = u0 % T3_DOWN_KS;
            Piece pc[1] = {Piece{p.act, 2 * T3_INTER, T3_INTER}};
            load_x<1>(xs, pc, 1, active, pp, bb, s * 1024 + kw);
            make_frags(A, xs);
            consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          }
          cbar();
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (gg < 2 * npos && (gg & 1) < p.B) {
              const int u = t3_plan_unit(PH_DOWN, c, G, j);
              p.down[(size_t)(u % T3_DOWN_KS) * HS + gg * T3_H + (u / T3_DOWN_KS) * 8 + nn] = red_sum(red, j, o);
            }
          }
        }
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
      }
      // ---------- phase F: final RMSNorm + speech_head (t3_hf_backend.py:100-103) ----------
      if (p.mode == 1 || it == n_iters - 1) {
        Piece pc[6];
        pc[0] = Piece{p.hres + ((L - 1) & 1) * HS, 2 * T3_H, T3_H};
        pc[1] = Piece{p.oproj, 2 * T3_H, T3_H};
        for (int s = 0; s < 4; ++s) pc[2 + s] = Piece{p.down + s * HS, 2 * T3_H, T3_H};
        load_x<6>(xs, pc, 6, active, pp, bb, kw);
        rmsnorm_rows(xs, p.final_norm, kw, ssq, warp, lane);
        make_frags(A, xs);
        consume_units(nhead, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
        cbar();
        const int plast = npos - 1;
        for (int idx = tid; idx < nhead * 64; idx += NCONS) {
          const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
          const int row = t3_plan_unit(PH_HEAD, c, G, j) * 8 + nn;
          if ((gg >> 1) == plast && (gg & 1) < p.B && row < T3_VOCAB)
            p.logits[(size_t)(gg & 1) * T3_VOCAB + row] = red_sum(red, j, o);
        }
        grid_barrier(p, bar_target, ctl);
        T3_PROF();
      }
      pos0 += npos;
      if (ctl->abort) break;
    }
    // publish state + tell the producer how far we got
    if (c == 0 && tid == 0) {
      p.st->pos = pos0;
      p.st->n_gen = n_gen;
      if (p.mode == 1 && p.n_out) *p.n_out = n_done;
    }
    cbar();
    if (tid == 0) {
      ctl->cons_k = rp.k;
      __threadfence_block();
      ctl->cons_done = 1;
    }
  }
  __syncthreads();
}

}  // namespace

// =================================================================================================
// host side of the C-ABI (include/cbx.h)
// =================================================================================================
struct cbx_t3 {
  CbxT3Config cfg;
  CbxT3Weights w;
  int G;
  T3Params base;        // pointers to owned scratch, filled at create
  void* slab;           // one allocation for all scratch
  int B;                // rows of the current utterance
  long long launches;
  unsigned long long* prof; int prof_it;
};

extern "C" int cbx_t3_grid(int device) {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    cbx_set_error("cbx_t3_grid: no CUDA device %d (this library has no CPU fallback)", device);
    return -1;
  }
  return prop.multiProcessorCount;
}
extern "C" int cbx_t3_plan_count(int ph, int c, int G) { return t3_plan_count(ph, c, G); }
extern "C" int cbx_t3_plan_unit(int ph, int c, int G, int j) { return t3_plan_unit(ph, c, G, j); }

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

extern "C" int cbx_t3_create(const CbxT3Config* cfg, const CbxT3Weights* w, cbx_t3** out) {
  CBX_CHECK(cfg && w && out, "cbx_t3_create: null argument");
  CBX_CHECK(cfg->n_layers >= 1 && cfg->n_layers <= 64, "cbx_t3_create: n_layers %d out of range", cfg->n_layers);
  CBX_CHECK(cfg->max_pos >= 8, "cbx_t3_create: max_pos too small");
  CBX_CUDA(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  CBX_CHECK(prop.major == 10, "cbx_t3_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  const int G = prop.multiProcessorCount;
  CBX_CHECK(G >= 129 && G * 0 + (1025 + G - 1) / G <= T3_MAXU, "cbx_t3_create: %d SMs unsupported by the plan", G);
  int coop = 0;
  CBX_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, cfg->device));
  CBX_CHECK(coop, "cbx_t3_create: device lacks cooperative launch");

  cbx_t3* h = new cbx_t3();
  h->cfg = *cfg; h->w = *w; h->G = G; h->B = 2; h->launches = 0; h->prof = nullptr; h->prof_it = -1;
  const size_t kvel = cfg->kv_fp32 ? 4 : 2;
  const size_t kv_bytes = (size_t)cfg->n_layers * 2 * T3_NH * cfg->max_pos * T3_HD * kvel;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t o_k = take(kv_bytes), o_v = take(kv_bytes);
  const size_t o_qkv = take(sizeof(float) * T3_MAX_NPOS * 2 * 3072);
  const size_t o_hres = take(sizeof(float) * 2 * T3_MAX_NPOS * 2 * T3_H);
  const size_t o_oproj = take(sizeof(float) * T3_MAX_NPOS * 2 * T3_H);
  const size_t o_act = take(sizeof(float) * T3_MAX_NPOS * 2 * T3_INTER);
  const size_t o_down = take(sizeof(float) * T3_DOWN_KS * T3_MAX_NPOS * 2 * T3_H);
  const size_t o_logits = take(sizeof(float) * 2 * T3_VOCAB);
  const size_t o_part = take(sizeof(float) * T3_MAX_NPOS * 32 * MAXS * PART_STRIDE);
  const size_t o_bar = take(256), o_st = take(256), o_seen = take(VOCAB_PAD);
  if (cudaMalloc(&h->slab, off) != cudaSuccess) {
    cbx_set_error("cbx_t3_create: cudaMalloc of %zu bytes failed", off);
    delete h;
    return CBX_ERR;
  }
  cudaMemset(h->slab, 0, off);
  uint8_t* s = static_cast<uint8_t*>(h->slab);
  T3Params& p = h->base;
  memset(&p, 0, sizeof(p));
  p.wstream = static_cast<const uint8_t*>(w->wstream);
  p.stride_bytes = (long long)w->stride_units * T3_UNIT_BYTES;
  p.ln1 = w->ln1; p.ln2 = w->ln2; p.final_norm = w->final_norm;
  p.speech_emb = w->speech_emb; p.speech_pos_emb = w->speech_pos_emb;
  p.rope_cos = w->rope_cos; p.rope_sin = w->rope_sin;
  p.n_layers = cfg->n_layers; p.max_pos = cfg->max_pos;
  p.kcache = s + o_k; p.vcache = s + o_v;
  p.qkv = reinterpret_cast<float*>(s + o_qkv); p.hres = reinterpret_cast<float*>(s + o_hres);
  p.oproj = reinterpret_cast<float*>(s + o_oproj); p.act = reinterpret_cast<float*>(s + o_act);
  p.down = reinterpret_cast<float*>(s + o_down); p.logits = reinterpret_cast<float*>(s + o_logits);
  p.part = reinterpret_cast<float*>(s + o_part);
  p.bar = reinterpret_cast<unsigned*>(s + o_bar); p.st = reinterpret_cast<T3State*>(s + o_st);
  p.seen = s + o_seen;
  CBX_CUDA(cudaFuncSetAttribute(t3_megakernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
  CBX_CUDA(cudaFuncSetAttribute(t3_megakernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
  int occ = 0;
  CBX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, t3_megakernel<__nv_bfloat16>, NTHREADS, SMEM_TOTAL));
  CBX_CHECK(occ >= 1, "cbx_t3_create: megakernel does not fit on an SM (smem %d B)", SMEM_TOTAL);
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_t3_destroy(cbx_t3* h) {
  if (!h) return CBX_OK;
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

static int t3_launch(cbx_t3* h, T3Params& p, cudaStream_t st) {
  CBX_CUDA(cudaMemsetAsync(p.bar, 0, sizeof(unsigned), st));
  void* args[] = {&p};
  const void* fn = h->cfg.kv_fp32 ? (const void*)t3_megakernel<float> : (const void*)t3_megakernel<__nv_bfloat16>;
  CBX_CUDA(cudaLaunchCooperativeKernel(fn, dim3(h->G), dim3(NTHREADS), args, SMEM_TOTAL, st));
  h->launches += 1;
  return CBX_OK;
}

extern "C" int cbx_t3_prefill(cbx_t3* h, const float* embeds, int B, int P, void* stream) {
  CBX_CHECK(h && embeds, "cbx_t3_prefill: null argument");
  CBX_CHECK(B == 1 || B == 2, "cbx_t3_prefill: B must be 1 or 2 (got %d)", B);
  CBX_CHECK(P >= 1 && P <= h->cfg.max_pos, "cbx_t3_prefill: P=%d exceeds the KV slab (max_pos %d)", P, h->cfg.max_pos);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_CUDA(cudaSetDevice(h->cfg.device));
  T3Params p = h->base;
  h->B = B;
  // reset per-utterance state: cache length 0, generated set = {BOS 6561} (t3.py:298)
  CBX_CUDA(cudaMemsetAsync(p.st, 0, sizeof(T3State), st));
  CBX_CUDA(cudaMemsetAsync(p.seen, 0, VOCAB_PAD, st));
  CBX_CUDA(cudaMemsetAsync(p.seen + 6561, 1, 1, st));
  p.mode = 0; p.B = B; p.embeds = embeds; p.P = P;
  return t3_launch(h, p, st);
}

extern "C" int cbx_t3_decode(cbx_t3* h, const CbxSampleParams* sp, int max_new_tokens, const int32_t* forced_tokens,
                             int32_t* tokens_out, int32_t* n_out, float* logits_out, void* stream) {
  CBX_CHECK(h && sp && tokens_out && n_out, "cbx_t3_decode: null argument");
  CBX_CHECK(max_new_tokens >= 1, "cbx_t3_decode: max_new_tokens must be >= 1");
  CBX_CHECK(sp->temperature > 0.f, "cbx_t3_decode: temperature must be > 0");
  CBX_CHECK(sp->repetition_penalty > 0.f, "cbx_t3_decode: repetition_penalty must be > 0");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_CUDA(cudaSetDevice(h->cfg.device));
  T3Params p = h->base;
  p.mode = 1; p.B = h->B; p.sp = *sp; p.max_new = max_new_tokens; p.forced = forced_tokens;
  p.tokens_out = tokens_out; p.n_out = n_out; p.logits_out = logits_out;
  p.prof = h->prof; p.prof_it = h->prof_it;
  return t3_launch(h, p, st);
}

extern "C" int cbx_t3_set_profile(cbx_t3* h, uint64_t* stamps, int it) {
  CBX_CHECK(h, "cbx_t3_set_profile: null handle");
  h->prof = reinterpret_cast<unsigned long long*>(stamps);
  h->prof_it = it;
  return CBX_OK;
}

extern "C" int cbx_t3_state(cbx_t3* h, int32_t* pos, int32_t* n_gen, int32_t* err, int64_t* launches) {
  CBX_CHECK(h, "cbx_t3_state: null handle");
  T3State s;
  CBX_CUDA(cudaMemcpy(&s, h->base.st, sizeof(s), cudaMemcpyDeviceToHost));
  if (pos) *pos = s.pos;
  if (n_gen) *n_gen = s.n_gen;
  if (err) *err = s.err;
  if (launches) *launches = h->launches;
  return CBX_OK;
}
