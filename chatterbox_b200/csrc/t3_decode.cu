// T3 (0.5B Llama) prefill + decode loop as ONE persistent cooperative kernel for sm_100a.
//
// Replaces the reference's per-token Python loop (src/chatterbox/models/t3/t3.py:319-367 and
// T3HuggingfaceBackend.forward, t3/inference/t3_hf_backend.py:73-116): ~600 library launches and a
// host sync per token become zero launches per token.  The step is HBM-bound (1.02 GB of bf16
// weights per token at batch 2), so the design is a weight-streaming pipeline:
//   * one CTA per SM (cooperative launch for co-residency), 8 consumer warps + 1 producer warp;
//   * the producer walks this CTA's pre-packed, contiguous weight stream with 16 KB TMA bulk
//     copies (cp.async.bulk -> UBLKCP) into a 9-deep shared-memory ring guarded by mbarriers.
//     Weights do not depend on activations, so the producer never waits on the consumers' data
//     dependencies: HBM keeps streaming while dependent phases hand activations over;
//   * consumers multiply each 8-row x 1024-k unit against the (<= 8) activation rows with
//     mma.sync m16n8k16 (weights are the B operand; activations are split x = hi + lo in bf16 and
//     occupy A rows 0-7 / 8-15, so products carry ~16 mantissa bits and accumulate in fp32);
//   * dependent phases (qkv -> attention -> combine -> o_proj -> gate/up -> down) hand their
//     activations over through "LL" buffers in L2: every fp32 value travels in one 8-byte word
//     together with a sequence flag ({value, flag}, one atomic 64-bit store), so a consumer polls
//     the data itself and there is NO grid barrier and NO fence on the critical path;
//   * residual adds, RMSNorm, RoPE, SwiGLU, the split-KV combine, CFG, logit processing, sampling,
//     EOS and the next-token embedding are fused into the phase prologues/epilogues; the residual
//     stream lives in registers (replicated per CTA), so nothing but weights and the KV slab
//     touches HBM.
// The same kernel runs chunked prefill (4 positions x 2 CFG rows per weight pass).
#include "common.cuh"
#include "t3_plan.h"
#include "../../include/cbx.h"

#include <cooperative_groups.h>
#include <math.h>
#include <stdarg.h>

namespace {

constexpr int NSLOT = 9;                  // ring depth (16 KB units)
constexpr int NCONS_WARPS = 8;
constexpr int NCONS = NCONS_WARPS * 32;   // consumer threads
constexpr int NTHREADS = NCONS + 32;      // + producer warp
constexpr int MAXS = 4;                   // max KV splits per (row, head) pair
constexpr int PART_STRIDE = 68;           // 64 o + m + l (+2 pad)
constexpr int STOP_SPEECH = 6562;
constexpr int VOCAB_PAD = 8200;
constexpr int MAXROWS = 2 * T3_MAX_NPOS;  // activation rows per weight pass: (position, CFG row)
constexpr int STAGE_LD = 132;             // padded row stride of the per-warp staging tile (floats)
constexpr float RMS_EPS = 1e-5f;

constexpr int SMEM_RING = NSLOT * T3_UNIT_BYTES;
constexpr int SMEM_BARS = SMEM_RING;            // 2*NSLOT mbarriers
constexpr int SMEM_CTL = SMEM_BARS + 256;       // control words
constexpr int SMEM_SCRATCH = SMEM_CTL + 256;    // phase scratch (union)
constexpr int SCRATCH_BYTES = VOCAB_PAD * 4 + 1024;
constexpr int SMEM_STAGE = SMEM_SCRATCH + SCRATCH_BYTES;              // [8 warps][MAXROWS][STAGE_LD] floats
constexpr int STAGE_BYTES = NCONS_WARPS * MAXROWS * STAGE_LD * 4;
constexpr int SMEM_SEEN = SMEM_STAGE + STAGE_BYTES;                   // generated-id set (CTA 0)
constexpr int SMEM_LN = SMEM_SEEN + 8208;                              // [2][1024] floats: prefetched norm gains
constexpr int SMEM_TOTAL = SMEM_LN + 2 * T3_H * 4;
static_assert(SMEM_TOTAL <= 232448, "shared memory budget");

// LL exchange buffers (elements are {fp32 bits, flag} 8-byte words), offsets in words
enum { LLP_QKV = 0, LLP_PART = 1, LLP_ATTN = 2, LLP_OPROJ = 3, LLP_ACT = 4, LLP_DOWN = 5, LLP_LOGITS = 6, LLP_EMB = 7 };
constexpr size_t LL_QKV = 0;
constexpr size_t LL_PART = LL_QKV + (size_t)MAXROWS * 3072;
constexpr size_t LL_ATTN = LL_PART + (size_t)T3_MAX_NPOS * 32 * MAXS * PART_STRIDE;
constexpr size_t LL_OPROJ = LL_ATTN + (size_t)MAXROWS * T3_H;
constexpr size_t LL_ACT = LL_OPROJ + (size_t)MAXROWS * T3_H;
constexpr size_t LL_DOWN = LL_ACT + (size_t)MAXROWS * T3_INTER;
constexpr size_t LL_LOGITS = LL_DOWN + (size_t)T3_DOWN_KS * MAXROWS * T3_H;
constexpr size_t LL_EMB = LL_LOGITS + (size_t)2 * VOCAB_PAD;
constexpr size_t LL_WORDS = LL_EMB + T3_H + 8;

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
  unsigned long long* ll;       // LL exchange words (zeroed before every launch)
  float* logits;                // [2][8194] plain copy of the last logits (crosses launches)
  T3State* st; unsigned char* seen;
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

// ---- LL words: {value, flag} in one 64-bit relaxed store / load --------------------------------
__device__ __forceinline__ unsigned ll_flag(int it, int L, int l, int ph) {
  return 1u + (unsigned)((it * (L + 1) + l) * 8 + ph);
}
__device__ __forceinline__ void ll_store(unsigned long long* p, float v, unsigned flag) {
  const unsigned long long w = ((unsigned long long)flag << 32) | (unsigned long long)__float_as_uint(v);
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long ll_ld1(const unsigned long long* p) {
  unsigned long long w;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return w;
}
__device__ __forceinline__ void ll_ld2(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ bool ll_ok(unsigned long long w, unsigned flag) { return (unsigned)(w >> 32) == flag; }
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((unsigned)w); }

// watchdog bookkeeping of a poll loop; returns true when the caller must give up
__device__ __forceinline__ bool poll_timeout(unsigned& spins, unsigned long long& t0, Ctl* ctl, T3State* st) {
  if (ctl->abort) return true;
  if ((++spins & 255u) == 0) {
    if (t0 == 0) t0 = gtime_ns();
    else if (gtime_ns() - t0 > WATCHDOG_NS || __ldcg(&st->err)) {
      ctl->abort = 1;
      atomicExch(&st->err, 1);
      return true;
    }
  }
  return false;
}

// poll n (<= N) scattered LL words; w[i] receives the word at ptr[i]
template <int N>
__device__ __forceinline__ void ll_poll(const unsigned long long* (&ptr)[N], int n, unsigned flag,
                                        unsigned long long (&w)[N], Ctl* ctl, T3State* st) {
  unsigned spins = 0; unsigned long long t0 = 0;
  for (;;) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < N; ++i) if (i < n) w[i] = ll_ld1(ptr[i]);
#pragma unroll
    for (int i = 0; i < N; ++i) if (i < n) ok = ok && ll_ok(w[i], flag);
    if (ok) return;
    if (poll_timeout(spins, t0, ctl, st)) return;
  }
}

struct Piece { const unsigned long long* ptr; size_t row_stride; unsigned flag; };   // row_stride in words

// Warp-cooperative load of this warp's 128-wide k slice of R activation rows from NP LL pieces
// (summed) into the warp's staging tile stage[row][STAGE_LD].  Row r = (position r>>1, CFG row r&1)
// is loaded only if r < nrows and (r&1) < B.  All lanes issue 2*R/PAR... loads per round and poll
// them together, so one exchange costs about one L2 round trip.
template <int R, int PAR>
__device__ __forceinline__ void ll_stage(float* stage, const Piece* pc, int np, int kbase, int nrows, int B, int lane,
                                         Ctl* ctl, T3State* st) {
  constexpr int T = 2 * R;      // v2 loads per lane per piece: R rows * 64 v2 / 32 lanes
  float acc[T][2];
#pragma unroll
  for (int t = 0; t < T; ++t) acc[t][0] = acc[t][1] = 0.f;
  for (int p0 = 0; p0 < np; p0 += PAR) {
    unsigned long long wa[PAR][T], wb[PAR][T];
    unsigned spins = 0; unsigned long long t0 = 0;
    for (;;) {
      bool ok = true;
#pragma unroll
      for (int q = 0; q < PAR; ++q) {
        if (p0 + q < np) {
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const int i = lane + 32 * t, row = i >> 6, c2 = i & 63;
            if (row < nrows && (row & 1) < B)
              ll_ld2(pc[p0 + q].ptr + (size_t)row * pc[p0 + q].row_stride + kbase + 2 * c2, wa[q][t], wb[q][t]);
          }
        }
      }
#pragma unroll
      for (int q = 0; q < PAR; ++q) {
        if (p0 + q < np) {
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const int row = (lane + 32 * t) >> 6;
            if (row < nrows && (row & 1) < B)
              ok = ok && ll_ok(wa[q][t], pc[p0 + q].flag) && ll_ok(wb[q][t], pc[p0 + q].flag);
          }
        }
      }
      if (ok) break;
      if (poll_timeout(spins, t0, ctl, st)) break;
    }
#pragma unroll
    for (int q = 0; q < PAR; ++q) {
      if (p0 + q < np) {
#pragma unroll
        for (int t = 0; t < T; ++t) {
          const int row = (lane + 32 * t) >> 6;
          if (row < nrows && (row & 1) < B) { acc[t][0] += ll_val(wa[q][t]); acc[t][1] += ll_val(wb[q][t]); }
        }
      }
    }
  }
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int i = lane + 32 * t, row = i >> 6, c2 = i & 63;
    *reinterpret_cast<float2*>(stage + row * STAGE_LD + 2 * c2) = make_float2(acc[t][0], acc[t][1]);
  }
  __syncwarp();
}

// thread (g = lane>>2, tig = lane&3) reads its fragment values: row g, k = jj*32 + tig*8 + e of the warp slice
__device__ __forceinline__ void stage_read(float (&xs)[4][8], const float* stage, int g, int tig) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    const float4 a = *reinterpret_cast<const float4*>(stage + g * STAGE_LD + jj * 32 + tig * 8);
    const float4 b = *reinterpret_cast<const float4*>(stage + g * STAGE_LD + jj * 32 + tig * 8 + 4);
    xs[jj][0] = a.x; xs[jj][1] = a.y; xs[jj][2] = a.z; xs[jj][3] = a.w;
    xs[jj][4] = b.x; xs[jj][5] = b.y; xs[jj][6] = b.z; xs[jj][7] = b.w;
  }
  __syncwarp();
}

// ---- activation rows -> A fragments -------------------------------------------------------------
// Lane (g = lane>>2, tig = lane&3) owns activation row g = (pos pp = g>>1, CFG row bb = g&1) and, for
// each of the warp's four k32-blocks jj, the 8 consecutive k at warp*128 + jj*32 + tig*8.
__device__ __forceinline__ void make_frags(uint32_t (&A)[4][8], const float (&xs)[4][8]) {
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    split_hi_lo(xs[jj][0], xs[jj][1], A[jj][0], A[jj][1]);   // step 0: a0 (hi), a1 (lo)  k = 2tig,2tig+1
    split_hi_lo(xs[jj][2], xs[jj][3], A[jj][2], A[jj][3]);   //         a2 (hi), a3 (lo)  k = 2tig+8,+9
    split_hi_lo(xs[jj][4], xs[jj][5], A[jj][4], A[jj][5]);   // step 1
    split_hi_lo(xs[jj][6], xs[jj][7], A[jj][6], A[jj][7]);
  }
}

// Norm gains travel through a double-buffered shared-memory prefetch (cp.async issued one phase
// ahead), so the RMSNorm never waits on L2/HBM for them.
__device__ __forceinline__ void ln_prefetch(float* lnbuf, int qq, const float* __restrict__ lnw, int tid) {
  const uint32_t dst = smem_u32(lnbuf + (qq & 1) * T3_H + tid * 4);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(lnw + tid * 4) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// RMSNorm over the full 1024-wide row (reference: HF LlamaRMSNorm, eps 1e-5): the row is spread
// over the 4 tig lanes of 8 warps; partial sums of squares meet in shared memory.  out = w * (x * rsqrt)
__device__ __forceinline__ void rmsnorm_rows(float (&out)[4][8], const float (&xs)[4][8], const float* lnbuf, int qq,
                                             int kw, float* ssq, int warp, int lane, unsigned long long* dbg = nullptr) {
  if (dbg) dbg[0] = gtime_ns();
  float s = 0.f;
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
#pragma unroll
    for (int e = 0; e < 8; ++e) s += xs[jj][e] * xs[jj][e];
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  const int g = lane >> 2;
  if ((lane & 3) == 0) ssq[warp * 8 + g] = s;
  if (dbg) dbg[1] = gtime_ns();
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (dbg) dbg[2] = gtime_ns();
  cbar();
  if (dbg) dbg[3] = gtime_ns();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) tot += ssq[w * 8 + g];
  const float scale = 1.0f / sqrtf(tot * (1.0f / T3_H) + RMS_EPS);
  const float* lw = lnbuf + (qq & 1) * T3_H + kw;
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    const float4 a = *reinterpret_cast<const float4*>(lw + jj * 32);
    const float4 b = *reinterpret_cast<const float4*>(lw + jj * 32 + 4);
    out[jj][0] = a.x * (xs[jj][0] * scale); out[jj][1] = a.y * (xs[jj][1] * scale);
    out[jj][2] = a.z * (xs[jj][2] * scale); out[jj][3] = a.w * (xs[jj][3] * scale);
    out[jj][4] = b.x * (xs[jj][4] * scale); out[jj][5] = b.y * (xs[jj][5] * scale);
    out[jj][6] = b.z * (xs[jj][6] * scale); out[jj][7] = b.w * (xs[jj][7] * scale);
  }
  if (dbg) dbg[4] = gtime_ns();
}

// ---- consume n weight units from the ring: red[j][warp][64] = this warp's k-slice partial of unit j
struct RingPos { int slot; uint32_t parity; int k; unsigned long long wait_ns, mma_ns; };

__device__ __forceinline__ void consume_units(int n, RingPos& rp, const uint32_t (&A)[4][8], float* red,
                                              const uint8_t* ring, uint64_t* full, uint64_t* empty, int warp, int lane,
                                              Ctl* ctl, T3State* st) {
  float2 prev = make_float2(0.f, 0.f);
  for (int j = 0; j < n; ++j) {
#ifdef T3_DEBUG_WAIT
    const unsigned long long tw0 = gtime_ns();
#endif
    mbar_wait(&full[rp.slot], rp.parity, ctl, st);
#ifdef T3_DEBUG_WAIT
    const unsigned long long tw1 = gtime_ns();
    rp.wait_ns += tw1 - tw0;
#endif
    const uint4* wp = reinterpret_cast<const uint4*>(ring + rp.slot * T3_UNIT_BYTES) + warp * 128 + lane;
    uint4 w[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) w[jj] = wp[jj * 32];
    // the previous unit's result is stored only now, so its MMA chain overlaps this unit's loads
    if (j > 0) reinterpret_cast<float2*>(red + ((j - 1) * NCONS_WARPS + warp) * 64)[lane] = prev;
    float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};   // two independent accumulator chains
    mma_bf16_16816(c0, A[0][0], A[0][1], A[0][2], A[0][3], w[0].x, w[0].y);
    mma_bf16_16816(c1, A[0][4], A[0][5], A[0][6], A[0][7], w[0].z, w[0].w);
    mma_bf16_16816(c0, A[1][0], A[1][1], A[1][2], A[1][3], w[1].x, w[1].y);
    mma_bf16_16816(c1, A[1][4], A[1][5], A[1][6], A[1][7], w[1].z, w[1].w);
    mma_bf16_16816(c0, A[2][0], A[2][1], A[2][2], A[2][3], w[2].x, w[2].y);
    mma_bf16_16816(c1, A[2][4], A[2][5], A[2][6], A[2][7], w[2].z, w[2].w);
    mma_bf16_16816(c0, A[3][0], A[3][1], A[3][2], A[3][3], w[3].x, w[3].y);
    mma_bf16_16816(c1, A[3][4], A[3][5], A[3][6], A[3][7], w[3].z, w[3].w);
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[rp.slot]);
    prev = make_float2((c0[0] + c0[2]) + (c1[0] + c1[2]), (c0[1] + c0[3]) + (c1[1] + c1[3]));
    if (++rp.slot == NSLOT) { rp.slot = 0; rp.parity ^= 1u; }
    ++rp.k;
#ifdef T3_DEBUG_WAIT
    rp.mma_ns += gtime_ns() - tw1;
#endif
  }
  if (n > 0) reinterpret_cast<float2*>(red + ((n - 1) * NCONS_WARPS + warp) * 64)[lane] = prev;
}

__device__ __forceinline__ float red_sum(const float* red, int j, int o) {
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) s += red[(j * NCONS_WARPS + w) * 64 + o];
  return s;
}

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

// ---- phase B: RoPE + KV append + split-KV attention over the slab + split combine ---------------
// (reference: HF LlamaAttention eager path: softmax(q k^T / 8 + causal) v in fp32; rotate_half RoPE)
// CTA c < S*npairs handles pair = c % npairs (CFG row b, head) and KV split c / npairs; it publishes
// {sum_i exp(s_i - m) v_i, m, l} through the LL part buffer, and the split-0 CTA of each pair merges
// the S partials into the attention output row segment (LL attn buffer).
template <typename KVT, int NPOS>
__device__ __noinline__ void attn_phase(const T3Params& p, int it, int layer, int pos0, int npos, float* scr, int warp,
                                        int lane, Ctl* ctl) {
  const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x, L = p.n_layers;
  const int npairs = p.B * T3_NH;
  int S = G / npairs; if (S > MAXS) S = MAXS;
  if (c >= S * npairs) return;
  const int pair = c % npairs, split = c / npairs, b = pair / T3_NH, head = pair % T3_NH;
  float* q_s = scr;                       // [4][64] (pre-scaled by 1/8)
  float* k_new = scr + 256;               // [4][64]
  float* v_new = scr + 512;               // [4][64]
  float* wpart = scr + 768;               // [8 warps][4][PART_STRIDE]
  KVT* kc = reinterpret_cast<KVT*>(p.kcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;
  KVT* vc = reinterpret_cast<KVT*>(p.vcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;

  float kr = 0.f, vr = 0.f;
  if (tid < npos * 64) {
    const int pp = tid >> 6, d = tid & 63;
    const unsigned long long* row = p.ll + LL_QKV + (size_t)(pp * 2 + b) * 3072 + head * 64;
    const unsigned long long* ptr[5] = {row + d, row + (d ^ 32), row + 1024 + d, row + 1024 + (d ^ 32),
                                              row + 2048 + d};
    unsigned long long w[5];
    ll_poll<5>(ptr, 5, ll_flag(it, L, layer, LLP_QKV), w, ctl, p.st);
    const float cs = __ldg(p.rope_cos + (size_t)(pos0 + pp) * 32 + (d & 31));
    const float sn = __ldg(p.rope_sin + (size_t)(pos0 + pp) * 32 + (d & 31));
    const float q = ll_val(w[0]), qo = ll_val(w[1]), k = ll_val(w[2]), ko = ll_val(w[3]), v = ll_val(w[4]);
    const float qr = q * cs + (d < 32 ? -qo : qo) * sn;
    kr = KV<KVT>::round(k * cs + (d < 32 ? -ko : ko) * sn);
    vr = KV<KVT>::round(v);
    q_s[pp * 64 + d] = qr * 0.125f;
    k_new[pp * 64 + d] = kr;
    v_new[pp * 64 + d] = vr;
  }
  cbar();

  const int lg = lane >> 3, sub = lane & 7;
  float qv[NPOS][8];
  Osm st[NPOS];
#pragma unroll
  for (int pp = 0; pp < NPOS; ++pp) {
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
    for (int pp = 0; pp < NPOS; ++pp) {
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
    for (int pp = 0; pp < NPOS; ++pp) {
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
  for (int pp = 0; pp < NPOS; ++pp) {
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
    float Ls = 0.f, O = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < NCONS_WARPS; ++w) {
        const float* wp = wpart + (w * T3_MAX_NPOS + pp) * PART_STRIDE;
        const float cw = (wp[64] == -INFINITY) ? 0.f : __expf(wp[64] - M);
        Ls += cw * wp[65];
        O += cw * wp[d];
      }
    }
    const unsigned fl = ll_flag(it, L, layer, LLP_PART);
    unsigned long long* out = p.ll + LL_PART + ((size_t)(pp * 32 + pair) * MAXS + split) * PART_STRIDE;
    ll_store(out + d, O, fl);
    if (d == 0) { ll_store(out + 64, M, fl); ll_store(out + 65, Ls, fl); }
    // KV append after the partial is on its way; the fence orders it before this CTA's later LL stores
    if (split == S - 1) {
      KV<KVT>::store(kc + (size_t)(pos0 + pp) * 64 + d, kr);
      KV<KVT>::store(vc + (size_t)(pos0 + pp) * 64 + d, vr);
      __threadfence();
    }
    if (split == 0) {
      // merge the S partials of this pair -> attention output segment
      const unsigned long long* const base = p.ll + LL_PART + ((size_t)(pp * 32 + pair) * MAXS) * PART_STRIDE;
      const unsigned long long* ptr[3 * MAXS];
#pragma unroll
      for (int s = 0; s < MAXS; ++s) {
        ptr[3 * s] = base + s * PART_STRIDE + d;
        ptr[3 * s + 1] = base + s * PART_STRIDE + 64;
        ptr[3 * s + 2] = base + s * PART_STRIDE + 65;
      }
      unsigned long long w[3 * MAXS];
      ll_poll<3 * MAXS>(ptr, 3 * S, fl, w, ctl, p.st);
      float Mx = -INFINITY;
#pragma unroll
      for (int s = 0; s < MAXS; ++s) if (s < S) Mx = fmaxf(Mx, ll_val(w[3 * s + 1]));
      float Lt = 0.f, Ot = 0.f;
#pragma unroll
      for (int s = 0; s < MAXS; ++s) {
        if (s < S) {
          const float ms = ll_val(w[3 * s + 1]);
          const float cw = (ms == -INFINITY) ? 0.f : __expf(ms - Mx);
          Lt += cw * ll_val(w[3 * s + 2]);
          Ot += cw * ll_val(w[3 * s]);
        }
      }
      ll_store(p.ll + LL_ATTN + (size_t)(pp * 2 + b) * T3_H + head * 64 + d, Ot / Lt, ll_flag(it, L, layer, LLP_ATTN));
    }
  }
}

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
__device__ __noinline__ void sample_phase(const T3Params& p, int it, int n_gen, float* sl, unsigned char* seen, Ctl* ctl, int warp,
                             int lane) {
  const int tid = threadIdx.x, L = p.n_layers;
  float* rs = sl + VOCAB_PAD;          // reduction scratch (64 floats)
  int* ri = reinterpret_cast<int*>(rs + 32);
  const CbxSampleParams& sp = p.sp;
  const bool cfg = sp.cfg_weight > 0.f && p.B == 2;
  // 1. CFG, temperature, repetition penalty; running max (first index on ties)
  float lmax = -INFINITY; int imax = 0x7fffffff;
  constexpr int NK = (T3_VOCAB + NCONS - 1) / NCONS;   // 33
  const unsigned lfl = ll_flag(it - 1, L, L, LLP_LOGITS);
  for (int k0 = 0; k0 < NK; k0 += 8) {
    float cv[8], uv[8];
    if (it == 0) {   // logits of the previous launch (prefill or an earlier decode call)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int i = tid + (k0 + k) * NCONS;
        cv[k] = (i < T3_VOCAB) ? __ldcg(p.logits + i) : 0.f;
        uv[k] = (i < T3_VOCAB && p.B == 2) ? __ldcg(p.logits + T3_VOCAB + i) : 0.f;
      }
    } else {
      unsigned long long wc[8], wu[8];
      unsigned spins = 0; unsigned long long t0 = 0;
      for (;;) {
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = tid + (k0 + k) * NCONS;
          if (i < T3_VOCAB) {
            wc[k] = ll_ld1(p.ll + LL_LOGITS + i);
            if (p.B == 2) wu[k] = ll_ld1(p.ll + LL_LOGITS + VOCAB_PAD + i);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = tid + (k0 + k) * NCONS;
          if (i < T3_VOCAB) ok = ok && ll_ok(wc[k], lfl) && (p.B != 2 || ll_ok(wu[k], lfl));
        }
        if (ok) break;
        if (poll_timeout(spins, t0, ctl, p.st)) break;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) { cv[k] = ll_val(wc[k]); uv[k] = (p.B == 2) ? ll_val(wu[k]) : 0.f; }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = tid + (k0 + k) * NCONS;
      if (i < T3_VOCAB) {
        const float c = cv[k], u = uv[k];
        if (p.logits_out) {
          p.logits_out[((size_t)it * p.B) * T3_VOCAB + i] = c;
          if (p.B == 2) p.logits_out[((size_t)it * p.B + 1) * T3_VOCAB + i] = u;
        }
        float x = cfg ? c + sp.cfg_weight * (c - u) : c;
        if (sp.temperature != 1.0f) x = x / sp.temperature;
        if (seen[i]) x = (x < 0.f) ? x * sp.repetition_penalty : x / sp.repetition_penalty;
        sl[i] = x;
        if (x > lmax) { lmax = x; imax = i; }
      }
    }
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
  const bool stop = sp.stop_on_eos && tok == STOP_SPEECH;
  if (tid == 0) {
    p.tokens_out[it] = tok;
    seen[tok] = 1;
    p.seen[tok] = 1;
    if (stop) p.st->stop = 1;
  }
  // 5. next-token embedding: speech_emb[tok] + speech_pos_emb[n_gen + 1]  (t3.py:351-356); both CFG rows get the
  //    same vector, published once together with the control word {token | stop bit}
  {
    const unsigned fl = ll_flag(it, L, 0, LLP_EMB);
    const float4 e = __ldg(reinterpret_cast<const float4*>(p.speech_emb + (size_t)tok * T3_H) + tid);
    const float4 q = __ldg(reinterpret_cast<const float4*>(p.speech_pos_emb + (size_t)(n_gen + 1) * T3_H) + tid);
    unsigned long long* dst = p.ll + LL_EMB + 4 * tid;
    ll_store(dst + 0, e.x + q.x, fl); ll_store(dst + 1, e.y + q.y, fl);
    ll_store(dst + 2, e.z + q.z, fl); ll_store(dst + 3, e.w + q.w, fl);
    if (tid == 0) ll_store(p.ll + LL_EMB + T3_H, __uint_as_float((unsigned)tok | (stop ? 0x80000000u : 0u)), fl);
  }
  cbar();   // seen[] update visible to the next call
}

// ---- the megakernel -----------------------------------------------------------------------------
template <typename KVT, int NPOS>
__global__ void __launch_bounds__(NTHREADS, 1) t3_megakernel(const T3Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SMEM_BARS);
  uint64_t* empty = full + NSLOT;
  Ctl* ctl = reinterpret_cast<Ctl*>(smem + SMEM_CTL);
  float* scr = reinterpret_cast<float*>(smem + SMEM_SCRATCH);
  float* red = scr;                       // [MAXU][8][64]
  float* ssq = scr + T3_MAXU * 8 * 64;    // [8][8]
  float* stage_all = reinterpret_cast<float*>(smem + SMEM_STAGE);
  unsigned char* seen = smem + SMEM_SEEN;
  float* lnbuf = reinterpret_cast<float*>(smem + SMEM_LN);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, c = blockIdx.x;
  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NCONS_WARPS); }
    ctl->abort = 0; ctl->cons_done = 0; ctl->cons_k = 0; ctl->tok = 0;
    mbar_fence_init();
  }
  for (int i = tid; i < NCONS_WARPS * MAXROWS * STAGE_LD; i += NTHREADS) stage_all[i] = 0.f;
  if (c == 0) for (int i = tid; i < VOCAB_PAD; i += NTHREADS) seen[i] = p.seen[i];
  __syncthreads();

  const int L = p.n_layers;
  const int upl = t3_plan_units_per_layer(c, G);
  const int nhead = t3_plan_count(PH_HEAD, c, G);
  const int n_iters = (p.mode == 0) ? (p.P + NPOS - 1) / NPOS : p.max_new;

  if (warp == NCONS_WARPS) {
    // ===== producer warp: stream this CTA's weight units through the ring =====
    const uint64_t policy = l2_policy_evict_first();
    const uint8_t* base = p.wstream + (size_t)c * p.stride_bytes;
    int slot = 0; uint32_t parity = 1; int k = 0;
    bool stopped = false;
    for (int it = 0; it < n_iters && !stopped; ++it) {
      const bool head = (p.mode == 1) || (it == n_iters - 1);
      const int n_units = L * upl + (head ? nhead : 0);
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
    constexpr int R = 2 * NPOS;
    constexpr int PAR = (NPOS == 1) ? 4 : 1;
    const int g = lane >> 2, tig = lane & 3, pp = g >> 1, bb = g & 1;
    const int kwb = warp * 128, kw = kwb + tig * 8;
    float* stage = stage_all + warp * MAXROWS * STAGE_LD;
    RingPos rp{0, 0u, 0, 0ull, 0ull};
    int pos0 = __ldcg(&p.st->pos);
    int n_gen = __ldcg(&p.st->n_gen);
    int n_done = 0;
    float xres[4][8];     // residual stream of row g, replicated in every CTA
    float xs[4][8];
    uint32_t A[4][8];
    int prof_k = 0;
    int qq = 0;           // running index of the norm whose gains sit in lnbuf[qq & 1]
    ln_prefetch(lnbuf, 0, p.ln1, tid);
#define T3_PROF() do { if (p.prof && c == 0 && tid == 0 && it == p.prof_it) p.prof[prof_k++] = gtime_ns(); } while (0)
#define T3_SUB(k) do { if (p.prof && c == 0 && tid == 0 && it == p.prof_it && l == 5) p.prof[200 + (k)] = gtime_ns(); } while (0)
#define T3_ROW_OK(gg) ((gg) < 2 * npos && ((gg) & 1) < p.B)

    for (int it = 0; it < n_iters; ++it) {
      int npos = 1;
      if (p.mode == 1) {
        T3_PROF();
        if (c == 0) sample_phase(p, it, n_gen, scr, seen, ctl, warp, lane);
        // control word of this iteration: token id | stop bit
        const unsigned fl = ll_flag(it, L, 0, LLP_EMB);
        unsigned long long w;
        unsigned spins = 0; unsigned long long t0 = 0;
        for (;;) {
          w = ll_ld1(p.ll + LL_EMB + T3_H);
          if (ll_ok(w, fl)) break;
          if (poll_timeout(spins, t0, ctl, p.st)) break;
        }
        ++n_gen; ++n_done;
        if (((unsigned)w >> 31) || ctl->abort) break;
        T3_PROF();
      } else {
        npos = min(NPOS, p.P - it * NPOS);
      }
      const int nrows = 2 * npos;
      const bool active = T3_ROW_OK(g);

      for (int l = 0; l < L; ++l) {
        // ---------- phase A: x_l = residual; RMSNorm; fused q|k|v GEMV ----------
        {
          T3_SUB(0);
          if (l == 0) {
            if (p.mode == 0) {
#pragma unroll
              for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                for (int e = 0; e < 8; ++e) xres[jj][e] = 0.f;
              if (active) {
                const float* src = p.embeds + ((size_t)bb * p.P + (size_t)it * NPOS + pp) * T3_H + kw;
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                  const float4 a = __ldg(reinterpret_cast<const float4*>(src + jj * 32));
                  const float4 b = __ldg(reinterpret_cast<const float4*>(src + jj * 32 + 4));
                  xres[jj][0] = a.x; xres[jj][1] = a.y; xres[jj][2] = a.z; xres[jj][3] = a.w;
                  xres[jj][4] = b.x; xres[jj][5] = b.y; xres[jj][6] = b.z; xres[jj][7] = b.w;
                }
              }
            } else {
              Piece pc[1] = {Piece{p.ll + LL_EMB, 0, ll_flag(it, L, 0, LLP_EMB)}};
              ll_stage<R, 1>(stage, pc, 1, kwb, nrows, p.B, lane, ctl, p.st);
              stage_read(xres, stage, g, tig);
            }
          } else {
            Piece pc[T3_DOWN_KS];
#pragma unroll
            for (int s = 0; s < T3_DOWN_KS; ++s)
              pc[s] = Piece{p.ll + LL_DOWN + (size_t)s * MAXROWS * T3_H, T3_H, ll_flag(it, L, l - 1, LLP_DOWN)};
            ll_stage<R, PAR>(stage, pc, T3_DOWN_KS, kwb, nrows, p.B, lane, ctl, p.st);
            stage_read(xs, stage, g, tig);
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
#pragma unroll
              for (int e = 0; e < 8; ++e) xres[jj][e] += xs[jj][e];
          }
          T3_SUB(1);
          rmsnorm_rows(xs, xres, lnbuf, qq, kw, ssq, warp, lane);
          ++qq;
          ln_prefetch(lnbuf, qq, p.ln2 + (size_t)l * T3_H, tid);
          make_frags(A, xs);
          T3_SUB(2);
          const int n = t3_plan_count(PH_QKV, c, G);
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          T3_SUB(3);
          cbar();
          T3_SUB(4);
          const unsigned fl = ll_flag(it, L, l, LLP_QKV);
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (T3_ROW_OK(gg))
              ll_store(p.ll + LL_QKV + (size_t)gg * 3072 + t3_plan_unit(PH_QKV, c, G, j) * 8 + nn, red_sum(red, j, o), fl);
          }
          T3_SUB(5);
          cbar();
          T3_SUB(6);
        }
        T3_PROF();
        // ---------- phase B: RoPE, KV append, split-KV attention, split combine ----------
        attn_phase<KVT, NPOS>(p, it, l, pos0, npos, scr, warp, lane, ctl);
        T3_PROF();
        // ---------- phase C: o_proj ----------
        {
          Piece pc[1] = {Piece{p.ll + LL_ATTN, T3_H, ll_flag(it, L, l, LLP_ATTN)}};
          ll_stage<R, 1>(stage, pc, 1, kwb, nrows, p.B, lane, ctl, p.st);
          stage_read(xs, stage, g, tig);
          make_frags(A, xs);
          const int n = t3_plan_count(PH_O, c, G);
          cbar();    // the attention phase used the scratch that `red` aliases
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          cbar();
          const unsigned fl = ll_flag(it, L, l, LLP_OPROJ);
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (T3_ROW_OK(gg))
              ll_store(p.ll + LL_OPROJ + (size_t)gg * T3_H + t3_plan_unit(PH_O, c, G, j) * 8 + nn, red_sum(red, j, o), fl);
          }
          cbar();
        }
        T3_PROF();
        // ---------- phase D: residual add; post-attention RMSNorm; gate/up GEMV; SwiGLU ----------
        {
          T3_SUB(8);
          Piece pc[1] = {Piece{p.ll + LL_OPROJ, T3_H, ll_flag(it, L, l, LLP_OPROJ)}};
          ll_stage<R, 1>(stage, pc, 1, kwb, nrows, p.B, lane, ctl, p.st);
          stage_read(xs, stage, g, tig);
          T3_SUB(9);
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
#pragma unroll
            for (int e = 0; e < 8; ++e) xres[jj][e] += xs[jj][e];
          rmsnorm_rows(xs, xres, lnbuf, qq, kw, ssq, warp, lane,
                       (p.prof && c == 0 && tid == 0 && it == p.prof_it && l == 5) ? p.prof + 216 : nullptr);
          ++qq;
          {
            const bool head_next = (l == L - 1) && (p.mode == 1 || it == n_iters - 1);
            ln_prefetch(lnbuf, qq, head_next ? p.final_norm : p.ln1 + (size_t)((l + 1) % L) * T3_H, tid);
          }
          make_frags(A, xs);
          T3_SUB(10);
          const int n = t3_plan_count(PH_GU, c, G);
          consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          T3_SUB(11);
          cbar();
          T3_SUB(12);
          const unsigned fl = ll_flag(it, L, l, LLP_ACT);
          for (int idx = tid; idx < n * 32; idx += NCONS) {
            const int j = idx >> 5, o2 = idx & 31, gg = o2 >> 2, nn = o2 & 3;
            if (T3_ROW_OK(gg)) {
              const float gate = red_sum(red, j, gg * 8 + nn), up = red_sum(red, j, gg * 8 + 4 + nn);
              const float silu = gate / (1.0f + __expf(-gate));
              ll_store(p.ll + LL_ACT + (size_t)gg * T3_INTER + t3_plan_unit(PH_GU, c, G, j) * 4 + nn, silu * up, fl);
            }
          }
          T3_SUB(13);
          cbar();
          T3_SUB(14);
        }
        T3_PROF();
        // ---------- phase E: down_proj (k-slice s of this CTA) ----------
        {
          const int n = t3_plan_count(PH_DOWN, c, G);
          if (n > 0) {
            const int s = t3_plan_unit(PH_DOWN, c, G, 0) % T3_DOWN_KS;
            Piece pc[1] = {Piece{p.ll + LL_ACT, T3_INTER, ll_flag(it, L, l, LLP_ACT)}};
            ll_stage<R, 1>(stage, pc, 1, s * 1024 + kwb, nrows, p.B, lane, ctl, p.st);
            stage_read(xs, stage, g, tig);
            make_frags(A, xs);
            consume_units(n, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
          }
          cbar();
          const unsigned fl = ll_flag(it, L, l, LLP_DOWN);
          for (int idx = tid; idx < n * 64; idx += NCONS) {
            const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
            if (T3_ROW_OK(gg)) {
              const int u = t3_plan_unit(PH_DOWN, c, G, j);
              ll_store(p.ll + LL_DOWN + (size_t)(u % T3_DOWN_KS) * MAXROWS * T3_H + (size_t)gg * T3_H +
                           (u / T3_DOWN_KS) * 8 + nn, red_sum(red, j, o), fl);
            }
          }
          cbar();
        }
        T3_PROF();
      }
      // ---------- phase F: final RMSNorm + speech_head (t3_hf_backend.py:100-103) ----------
      if (p.mode == 1 || it == n_iters - 1) {
        Piece pc[T3_DOWN_KS];
#pragma unroll
        for (int s = 0; s < T3_DOWN_KS; ++s)
          pc[s] = Piece{p.ll + LL_DOWN + (size_t)s * MAXROWS * T3_H, T3_H, ll_flag(it, L, L - 1, LLP_DOWN)};
        ll_stage<R, PAR>(stage, pc, T3_DOWN_KS, kwb, nrows, p.B, lane, ctl, p.st);
        stage_read(xs, stage, g, tig);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj)
#pragma unroll
          for (int e = 0; e < 8; ++e) xres[jj][e] += xs[jj][e];
        rmsnorm_rows(xs, xres, lnbuf, qq, kw, ssq, warp, lane);
        ++qq;
        ln_prefetch(lnbuf, qq, p.ln1, tid);
        make_frags(A, xs);
        consume_units(nhead, rp, A, red, ring, full, empty, warp, lane, ctl, p.st);
        cbar();
        const int plast = npos - 1;
        const unsigned fl = ll_flag(it, L, L, LLP_LOGITS);
        for (int idx = tid; idx < nhead * 64; idx += NCONS) {
          const int j = idx >> 6, o = idx & 63, gg = o >> 3, nn = o & 7;
          const int row = t3_plan_unit(PH_HEAD, c, G, j) * 8 + nn;
          if ((gg >> 1) == plast && (gg & 1) < p.B && row < T3_VOCAB) {
            const float v = red_sum(red, j, o);
            p.logits[(size_t)(gg & 1) * T3_VOCAB + row] = v;     // plain copy: read by the next launch
            if (p.mode == 1) ll_store(p.ll + LL_LOGITS + (size_t)(gg & 1) * VOCAB_PAD + row, v, fl);
          }
        }
        cbar();
        T3_PROF();
      }
#ifdef T3_DEBUG_WAIT
      if (p.prof && c == 0 && tid == 0 && it == p.prof_it) { p.prof[230] = rp.wait_ns; p.prof[231] = rp.mma_ns; p.prof[232] = rp.k; }
      rp.wait_ns = 0; rp.mma_ns = 0;
#endif
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

template <typename F>
static int t3_prepare_kernel(F fn) {
  CBX_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
  int occ = 0;
  CBX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, NTHREADS, SMEM_TOTAL));
  CBX_CHECK(occ >= 1, "cbx_t3_create: megakernel does not fit on an SM (smem %d B)", SMEM_TOTAL);
  return CBX_OK;
}

extern "C" int cbx_t3_create(const CbxT3Config* cfg, const CbxT3Weights* w, cbx_t3** out) {
  CBX_CHECK(cfg && w && out, "cbx_t3_create: null argument");
  CBX_CHECK(cfg->n_layers >= 1 && cfg->n_layers <= 64, "cbx_t3_create: n_layers %d out of range", cfg->n_layers);
  CBX_CHECK(cfg->max_pos >= 8, "cbx_t3_create: max_pos too small");
  CBX_CUDA(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  CBX_CHECK(prop.major == 10, "cbx_t3_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  const int G = prop.multiProcessorCount;
  CBX_CHECK(G >= 129 && (1025 + G - 1) / G <= T3_MAXU, "cbx_t3_create: %d SMs unsupported by the plan", G);
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
  const size_t o_ll = take(LL_WORDS * 8);
  const size_t o_logits = take(sizeof(float) * 2 * T3_VOCAB);
  const size_t o_st = take(256), o_seen = take(VOCAB_PAD);
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
  p.ll = reinterpret_cast<unsigned long long*>(s + o_ll);
  p.logits = reinterpret_cast<float*>(s + o_logits);
  p.st = reinterpret_cast<T3State*>(s + o_st);
  p.seen = s + o_seen;
  int rc = t3_prepare_kernel(t3_megakernel<__nv_bfloat16, 1>);
  if (!rc) rc = t3_prepare_kernel(t3_megakernel<__nv_bfloat16, T3_MAX_NPOS>);
  if (!rc) rc = t3_prepare_kernel(t3_megakernel<float, 1>);
  if (!rc) rc = t3_prepare_kernel(t3_megakernel<float, T3_MAX_NPOS>);
  if (rc) { cudaFree(h->slab); delete h; return CBX_ERR; }
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
  CBX_CUDA(cudaMemsetAsync(p.ll, 0, LL_WORDS * 8, st));     // flags restart at 1 in every launch
  void* args[] = {&p};
  const void* fn;
  if (p.mode == 0) fn = h->cfg.kv_fp32 ? (const void*)t3_megakernel<float, T3_MAX_NPOS>
                                       : (const void*)t3_megakernel<__nv_bfloat16, T3_MAX_NPOS>;
  else fn = h->cfg.kv_fp32 ? (const void*)t3_megakernel<float, 1> : (const void*)t3_megakernel<__nv_bfloat16, 1>;
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
