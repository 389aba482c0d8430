// T3 (0.5B Llama) prefill + decode loop as ONE persistent cooperative kernel for sm_100a.
//
// Replaces the reference's per-token Python loop (src/chatterbox/models/t3/t3.py:319-367 and
// T3HuggingfaceBackend.forward, t3/inference/t3_hf_backend.py:73-116): ~600 library launches and a
// host sync per token become zero launches per token.  The step is HBM-bound (1.02 GB of bf16
// weights per token at batch 2), so the design is a weight-streaming pipeline:
//   * one CTA per SM (cooperative launch for co-residency), 8 warps = 2 per scheduler, so every
//     thread may use up to 255 registers;
//   * this CTA's pre-packed, contiguous weight stream is consumed in 32 KB "units" (16 rows x 1024 k in MMA
//     fragment order).  Default build: the units of a step are loaded with coalesced 128-bit loads straight from
//     L2 into registers BEFORE the wait for the step's inputs (weights do not depend on activations), and
//     cp.async.bulk.prefetch.L2 pulls the units 6 ahead from HBM, so HBM streams ahead of the dependency chain.
//     -DT3_RING: the first design, 32 KB TMA bulk copies (cp.async.bulk) through a 4-deep shared-memory ring
//     with mbarrier completion (kept for A/B: 406 against 357 us/token);
//   * consumers multiply each unit against the CFG pair of activation rows with mma.sync m16n8k16: the WEIGHTS
//     are the A operand (16 rows per instruction, one 128-bit load feeds one MMA), the activations the B operand
//     with columns {row0 hi, row0 lo, row1 hi, row1 lo}: x = hi + lo in bf16, so products carry ~16 mantissa
//     bits and accumulate in fp32.  The 8 warps split K; partials meet in smem;
//   * dependent phases (qkv -> attention -> o_proj -> gate/up -> down) hand their activations
//     over through "LL" buffers in L2: every fp32 value travels in one 8-byte word together with
//     a sequence flag ({value, flag}, one atomic 64-bit store), so a consumer polls the data
//     itself and there is NO grid barrier and NO fence on the critical path;
//   * a lane owns four PAIRS of k, 32 apart, of one activation row, so every poll instruction of a warp is
//     coalesced (4 cache lines, not 16 LSU wavefronts); the owning lane splits its values into packed hi/lo bf16
//     words and a fragment-order tile turns them into B fragments with four 128-bit loads.  The residual slice and
//     the next norm gains live in shared memory (no spills).  The RMSNorm scale is a per-row scalar, so it is
//     applied to the OUTPUTS in the phase epilogue: the MMAs start without waiting for the sum of squares;
//   * attention: the split's K/V rows are copied to shared memory before the wait for q/k/v; contexts up to 544
//     positions per split take a two-pass tile path (all scores at once, per-warp redundant softmax statistics,
//     reduce-scatter output), longer ones and the fp32 slab an online-softmax stream with a cp.async ring;
//   * RoPE, SwiGLU, CFG, logit processing, sampling, EOS and the next-token embedding are fused
//     into the phase prologues/epilogues, so nothing but weights and the KV slab touches HBM.
// Prefill runs the same kernel one position per weight pass (mode 0).
#include "common.cuh"
#include "t3_plan.h"
#include "t3_prefill.h"
#include "../../include/cbx.h"

#include <cooperative_groups.h>
#include <math.h>
#include <stdlib.h>
#include <stdarg.h>
#include <vector>

namespace {

#ifndef T3_NSLOT
#define T3_NSLOT 4                        // 5 measured 1 % slower (the L1 left for the stack shrinks to ~4 KB)
#endif
constexpr int NSLOT = T3_NSLOT;           // ring depth (32 KB units)
#ifndef T3_SAMPLE_KB
#define T3_SAMPLE_KB 11
#endif
#if !defined(T3_RING)
#define T3_DIRECT 1                        // weight units straight from L2 into registers (-DT3_RING: the shared-memory TMA ring)
#endif
#ifndef T3_PF_DIST
#define T3_PF_DIST 6                      // L2 prefetch distance in units (0 = off)
#endif
constexpr int NCONS_WARPS = T3_KW;
constexpr int VPL = 1024 / NCONS_WARPS / 16;   // consecutive k one lane owns (8 or 4) = k16 steps of a warp's k-slice
static_assert(VPL == 8 || VPL == 4, "T3_KW must be 8 or 16");
constexpr int NCONS = NCONS_WARPS * 32;   // consumer threads
constexpr int NTHREADS = NCONS;
#define T3_STR_(x) #x
#define T3_STR(x) T3_STR_(x)
#ifndef T3_ATT_UNROLL
#define T3_ATT_UNROLL 1                  // unroll of the tile attention's row loops; 2 / 4 measured: 365.4 / 366.7 vs 364.6 us/token
#endif
#ifndef T3_MAXS
#define T3_MAXS 4
#endif
constexpr int MAXS = T3_MAXS;             // KV splits per (row, head) pair (4: 128 attention CTAs; 2: 64, half the partial words)
constexpr int PART_STRIDE = 68;           // 64 o + m + l (+2 pad)
constexpr int STOP_SPEECH = 6562;
constexpr int VOCAB_PAD = T3_VOCAB_PAD;
constexpr float RMS_EPS = 1e-5f;

#ifdef T3_DIRECT
constexpr int SMEM_RING = 0;                    // no weight ring: units go from L2 to registers
#else
constexpr int SMEM_RING = NSLOT * T3_UNIT_BYTES;
#endif
constexpr int SMEM_BARS = SMEM_RING;            // 2*NSLOT mbarriers
constexpr int SMEM_CTL = SMEM_BARS + 256;       // control words
constexpr int SMEM_RED = SMEM_CTL + 256;        // [2 phase parities][MAXU][warps][32] k-slice partials
constexpr int RED_FLOATS = T3_MAXU * NCONS_WARPS * 32;
constexpr int SMEM_SSQ = SMEM_RED + 2 * RED_FLOATS * 4;      // [2][warps][2 rows] partial sums of squares
constexpr int SMEM_XT = SMEM_SSQ + 256;                      // [warps][XT_WORDS] packed hi/lo activation fragments
constexpr int SMEM_SCRATCH = SMEM_XT + (NCONS_WARPS * (2 * (2 * 4 * (2 * VPL + 4) + 4)) * 4 + 15) / 16 * 16;   // fragment tiles (XT_WORDS per warp); then attention scratch / sampling logits (union)
// past positions an attention tile holds (+ 1 row for the position itself).  The direct build has the ring's 128 KB
// to spare: 544 positions per split cover contexts up to 2 176 (config 5) on the tile path.
#ifdef T3_DIRECT
constexpr int ATT_TP = 544;
#else
constexpr int ATT_TP = 96;
#endif
constexpr int ATT_TILE_BYTES = (64 + (ATT_TP + 1 + 31) / 32 * 32) * 4 + 2 * (ATT_TP + 1) * 72 * 2;
constexpr int SCRATCH_BYTES = (VOCAB_PAD * 4 + 1024 > ATT_TILE_BYTES ? VOCAB_PAD * 4 + 1024 : ATT_TILE_BYTES + 15) / 16 * 16;
constexpr int SMEM_SEEN = SMEM_SCRATCH + SCRATCH_BYTES;      // generated-id set (CTA 0), one bit per id
constexpr int SEEN_WORDS = (VOCAB_PAD + 31) / 32;
constexpr int SEEN_BYTES = (SEEN_WORDS * 4 + 15) / 16 * 16;
// copy of the kernel parameters for the out-of-line phases (attention, sampling): a reference to the parameter
// block itself would make the compiler keep it on the stack and turn every p.field in the step loop into a local
// load (L1 is small here: a miss costs an L2 round trip on the critical path)
constexpr int SMEM_PARAMS = SMEM_SEEN + SEEN_BYTES;
// norm gains of the NEXT RMSNorm, copied a phase ahead by cp.async (warp w copies exactly the 128 gains its own lanes
// read, so a warp-level sync is enough).  In registers they were 8 long-lived values the compiler spilled every step.
constexpr int SMEM_GN = SMEM_PARAMS + 512;
// residual stream (this CTA's replica, [warps][2 rows][16 VPL] like the fragment tile): 8 values per lane that live
// for a whole token and are touched in 3 of 5 steps — in registers the compiler spilled them to local memory
constexpr int SMEM_XR = SMEM_GN + T3_H * 4;
constexpr int SMEM_TOTAL = SMEM_XR + NCONS_WARPS * 2 * 16 * VPL * 4;
static_assert(NSLOT > 4 || SMEM_TOTAL + 1024 <= 196 * 1024, "4-slot build: stay inside the 196 KB carve-out (the rest of the 256 KB is L1)");
static_assert(SMEM_TOTAL <= 232448, "shared memory budget");
static_assert(2 * NCONS_WARPS * 2 * 4 <= 128, "ssq buffer: the sums use the first 128 B, the analyzer state the rest");
static_assert(SMEM_SCRATCH % 16 == 0 && SMEM_XT % 16 == 0, "alignment");

// LL exchange buffers (elements are {fp32 bits, flag} 8-byte words), offsets in words
enum { LLP_QKV = 0, LLP_PART = 1, LLP_ALIGN = 2, LLP_OPROJ = 3, LLP_ACT = 4, LLP_DOWN = 5, LLP_LOGITS = 6, LLP_EMB = 7 };
constexpr int AL_CAP = 2112;              // text columns the alignment analyzer can track (max_text_tokens 2048 + SOT / EOT, padded)
constexpr size_t LL_QKV = 0;                                          // [2][3072]
constexpr size_t LL_PART = LL_QKV + (size_t)2 * 3072;                 // [32 pairs][MAXS][PART_STRIDE]
constexpr size_t LL_OPROJ = LL_PART + (size_t)32 * MAXS * PART_STRIDE;   // [2][1024]
constexpr size_t LL_ACT = LL_OPROJ + (size_t)2 * T3_H;                // [2][4096]
constexpr size_t LL_DOWN = LL_ACT + (size_t)2 * T3_INTER;             // [4 k-slices][2][1024]
constexpr size_t LL_LOGITS = LL_DOWN + (size_t)T3_DOWN_KS * 2 * T3_H; // [2][VOCAB_PAD]
constexpr size_t LL_EMB = LL_LOGITS + (size_t)2 * VOCAB_PAD;          // [1024] + control word
// alignment analyzer (alignment_stream_analyzer.py): un-normalised softmax terms of the text columns, per head, and the
// (m, l) of every (head, KV split) of the analysed layer's conditional row
constexpr size_t LL_AL = LL_EMB + T3_H + 8;                           // [16 heads][AL_CAP]
constexpr size_t LL_ALML = LL_AL + (size_t)T3_NH * AL_CAP;            // [16][MAXS][2]
constexpr size_t LL_WORDS = LL_ALML + (size_t)T3_NH * MAXS * 2;
static_assert((AL_CAP + 64 + T3_NH * MAXS * 2 + T3_NH * 2) * 4 <= SCRATCH_BYTES, "analyzer scratch fits the sampling scratch");

struct T3State {   // lives in global memory, persists between launches of one utterance
  int pos;         // KV cache length
  int n_gen;       // speech tokens generated so far (next learned speech position = n_gen + 1)
  int stop;        // EOS seen
  int err;         // device-side error (1 = watchdog timeout)
};

// O(1) state of AlignmentStreamAnalyzer.step (alignment_stream_analyzer.py:89-154; oracle/t3_oracle.py AlignmentOracle):
// everything the reference reads off its growing alignment matrix A is a running statistic
struct AlState {
  int frame, text_pos, started, complete, T;
  float last2_prev, last2_cur;      // max of the last two columns of the two most recent rows      (A[-2:, -2:].max())
  float first4;                     // running max over the first four columns                      (A[:, :4].max())
  float tail0, tail1, tail2;        // column sums of the last three columns after completion       (A[completed_at:, -3:].sum(0))
  float rep;                        // sum of row maxima over columns [:S-5] after completion       (A[completed_at:, :-5].max(1).sum())
  int force, suppress, pad0, pad1;  // decision of the current step
};
static_assert(sizeof(AlState) == 64, "AlState layout");

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
  // alignment analyzer (off: al_layer < 0): analysed layer, first text position, text columns, query rows of the prompt
  int al_layer, al_i, al_n, al_rows0;
  const float* al_prefill;      // [al_rows0][AL_CAP] rows of the prompt prefill
  const float* al_inject;       // optional [rows][al_n]: rows used INSTEAD of the measured attention (tests of the state machine)
  float* al_rows_out;           // optional [rows][al_n]: every row the analyzer consumed, before the monotonic masking
  int* al_out;                  // optional [max_new][4]: text position, flags (started | complete << 1 | forced << 2 | suppressed << 3), argmax
  AlState* al_state;
};

struct Ctl {            // CTA-local control block in shared memory
  volatile int abort;   // set after a watchdog timeout: every later wait returns at once
  // cursor of the weight stream (next unit to copy): touched only by the one thread re-arming a slot
  int next_idx;         // unit index inside the per-CTA stream
  int next_iter;        // iteration that unit belongs to
  int plan_n[PH_COUNT];            // units of each phase on this CTA
  int plan_u[PH_COUNT][T3_MAXU];   // their unit ids
};
static_assert(sizeof(Ctl) <= 256, "control block");
static_assert(T3_MAXU <= NSLOT && T3_MAXU < NCONS_WARPS, "a step's units fit the ring; the last warp is free in the epilogue");
static_assert(sizeof(T3Params) <= 512 && SMEM_PARAMS % 16 == 0, "parameter copy");

using namespace cbx;

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long WATCHDOG_NS = 4000000000ull;

__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory"); }

// Wait on an mbarrier phase with a watchdog; returns false if aborted.  The slow path stays out of line: the
// layer loop has to stay resident in the instruction cache.
__device__ __noinline__ bool mbar_wait_slow(uint64_t* bar, uint32_t parity, Ctl* ctl, T3State* st) {
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
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, Ctl* ctl, T3State* st) {
  if (mbar_try_wait(bar, parity)) return true;
  return mbar_wait_slow(bar, parity, ctl, st);
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
// same with a compile-time byte offset folded into the address operand: polling many words off one base costs
// one pointer register pair, not one per word
template <int OFF>
__device__ __forceinline__ void ll_ld2_off(const unsigned long long* p, unsigned long long& a, unsigned long long& b) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2+%3];" : "=l"(a), "=l"(b) : "l"(p), "n"(OFF) : "memory");
}
#ifdef T3_ABL_NOPOLL   // ablation build (timing only, results are garbage): every word counts as arrived
__device__ __forceinline__ bool ll_ok(unsigned long long w, unsigned flag) { return true; }
#else
__device__ __forceinline__ bool ll_ok(unsigned long long w, unsigned flag) { return (unsigned)(w >> 32) == flag; }
#endif
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((unsigned)w); }

// watchdog bookkeeping of a poll loop; returns true when the caller must give up
__device__ __noinline__ bool poll_timeout_slow(unsigned long long& t0, Ctl* ctl, T3State* st) {
  if (t0 == 0) t0 = gtime_ns();
  else if (gtime_ns() - t0 > WATCHDOG_NS || __ldcg(&st->err)) {
    ctl->abort = 1;
    atomicExch(&st->err, 1);
    return true;
  }
  return false;
}
__device__ __forceinline__ bool poll_timeout(unsigned& spins, unsigned long long& t0, Ctl* ctl, T3State* st) {
  if (ctl->abort) return true;
  if ((++spins & 255u) == 0) return poll_timeout_slow(t0, ctl, st);
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

// ---- activation slices ---------------------------------------------------------------------------
// Warp w owns k in [128w, 128w + 128) of the two activation rows (CFG cond / uncond); lane owns row
// lane >> 4 and the VPL/2 PAIRS of k at 128w + 2 (lane & 15) + 32 r (r = 0..VPL/2-1): element e of its slice is
// k = kq + 32 (e >> 1) + (e & 1).  Consecutive lanes read consecutive 16-byte LL packets, so one warp-wide poll
// instruction touches 4 cache lines, not 16: the LSU issues one wavefront per line, and with 8 consecutive words
// per lane a poll round of the 8 warps cost ~1 us in wavefronts alone (measured by ablation).

// x[e] = sum over np (<= 4) pieces, `stride` words apart, of this lane's LL words (base already points at its first
// pair; fixed order -> deterministic sum).  One instance serves every phase: all loads of a round are in flight together.
template <int NP>
__device__ __forceinline__ void ll_load8_n(float (&x)[VPL], const unsigned long long* base, size_t stride, unsigned flag,
                                           bool active, Ctl* ctl, T3State* st) {
  unsigned long long w[NP][VPL];
  if (active) {
    unsigned spins = 0; unsigned long long t0 = 0;
    for (;;) {
      bool ok = true;
#pragma unroll
      for (int q = 0; q < NP; ++q) {
        const unsigned long long* pq = base + q * stride;
        ll_ld2_off<0>(pq, w[q][0], w[q][1]);
        ll_ld2_off<256>(pq, w[q][2], w[q][3]);
        if constexpr (VPL == 8) {
          ll_ld2_off<512>(pq, w[q][VPL - 4], w[q][VPL - 3]);
          ll_ld2_off<768>(pq, w[q][VPL - 2], w[q][VPL - 1]);
        }
      }
#pragma unroll
      for (int q = 0; q < NP; ++q) {
#pragma unroll
        for (int e = 0; e < VPL; ++e) ok = ok && ll_ok(w[q][e], flag);
      }
      if (ok) break;
      if (poll_timeout(spins, t0, ctl, st)) break;
    }
#pragma unroll
    for (int e = 0; e < VPL; ++e) {
      float sum = ll_val(w[0][e]);
#pragma unroll
      for (int q = 1; q < NP; ++q) sum += ll_val(w[q][e]);
      x[e] = sum;
    }
  } else {
#pragma unroll
    for (int e = 0; e < VPL; ++e) x[e] = 0.f;
  }
}
// the two piece counts the kernel uses get their own straight-line code (no per-word selects on np)
__device__ __forceinline__ void ll_load8(float (&x)[VPL], const unsigned long long* base, size_t stride, int np, unsigned flag,
                                         bool active, Ctl* ctl, T3State* st) {
#ifdef T3_ABL_NOLOAD
#pragma unroll
  for (int e = 0; e < VPL; ++e) x[e] = 0.001f * e;
  return;
#endif
  if (np == 1) ll_load8_n<1>(x, base, stride, flag, active, ctl, st);
  else ll_load8_n<T3_DOWN_KS>(x, base, stride, flag, active, ctl, st);
}

// B fragments of the warp's k-slice: Bf[s][0..1] for k16 step s.  Column n = lane >> 2 of the MMA is
// {row0 hi, row0 lo, row1 hi, row1 lo, 0, 0, 0, 0}.  The OWNING lane splits its pairs into packed bf16x2 hi / lo
// words (all 32 lanes busy, 6 instructions per pair) and scatters them into a per-warp tile laid out in fragment
// order [plane hi|lo][row][tig][k16 step][half]; lane (g < 4, tig) then fetches its 16 fragment words with four
// 128-bit loads and no arithmetic.  (Splitting after the transpose cost 212 instructions per step on the 16 reading
// lanes, 22 % of all instructions of the kernel.)  Strides are padded so the scattered stores are conflict-free.
constexpr int XT_T = 2 * VPL + 4;            // words per tig
constexpr int XT_R = 4 * XT_T;               // per row
constexpr int XT_P = 2 * XT_R + 4;           // per plane
constexpr int XT_WORDS = 2 * XT_P;           // per warp
static_assert(XT_T % 4 == 0 && XT_P % 4 == 0 && XT_WORDS % 4 == 0, "128-bit fragment loads");
__device__ __forceinline__ void make_bfrags(uint32_t (&Bf)[VPL][2], const float (&x)[VPL], uint32_t* xt, int lane) {
#ifdef T3_ABL_NOFRAGS
#pragma unroll
  for (int s = 0; s < VPL; ++s) { Bf[s][0] = __float_as_uint(x[s]); Bf[s][1] = 0u; }
  return;
#endif
  __syncwarp();     // previous readers of the tile are done
  {
    const int j = lane & 15;
    // k_local = 2 j + 32 r: k16 step 2 r + (j >> 3), tig = j & 3, half = (j >> 2) & 1
    uint32_t* dst = xt + (lane >> 4) * XT_R + (j & 3) * XT_T + (j >> 3) * 2 + ((j >> 2) & 1);
#pragma unroll
    for (int r = 0; r < VPL / 2; ++r) {
      uint32_t hi, lo;
      split_hi_lo(x[2 * r], x[2 * r + 1], hi, lo);
      dst[4 * r] = hi;
      dst[XT_P + 4 * r] = lo;
    }
  }
  __syncwarp();
  const int g = lane >> 2, tig = lane & 3;
#pragma unroll
  for (int s = 0; s < VPL; ++s) { Bf[s][0] = 0u; Bf[s][1] = 0u; }
  if (g < 4) {
    const uint4* src = reinterpret_cast<const uint4*>(xt + (g & 1) * XT_P + (g >> 1) * XT_R + tig * XT_T);
#pragma unroll
    for (int i = 0; i < VPL / 2; ++i) {
      const uint4 v = src[i];
      Bf[2 * i][0] = v.x; Bf[2 * i][1] = v.y; Bf[2 * i + 1][0] = v.z; Bf[2 * i + 1][1] = v.w;
    }
  }
}

// sum of squares of the lane's 8 values, reduced over the 16 lanes of the row; lane & 15 == 0 stores it
__device__ __forceinline__ void ssq_partial(const float (&x)[VPL], float* ssq, int warp, int lane) {
  float s = 0.f;
#pragma unroll
  for (int e = 0; e < VPL; ++e) s += x[e] * x[e];
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  s += __shfl_xor_sync(0xffffffffu, s, 8);
  if ((lane & 15) == 0) ssq[warp * 2 + (lane >> 4)] = s;
}
__device__ __forceinline__ float rms_scale(const float* ssq, int b) {
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) tot += ssq[w * 2 + b];
  return rsqrtf(tot * (1.0f / T3_H) + RMS_EPS);
}
// norm gains (or any fp32 vector) at this lane's k pairs, fetched one phase ahead into registers
__device__ __forceinline__ void ln_fetch(float (&g)[VPL], const float* __restrict__ lnw, int k) {
#pragma unroll
  for (int r = 0; r < VPL / 2; ++r) {
    const float2 a = __ldcg(reinterpret_cast<const float2*>(lnw + k + 32 * r));      // L2 only: keep the small L1 for the stack
    g[2 * r] = a.x; g[2 * r + 1] = a.y;
  }
}

// copy the 128 gains of this warp's k-slice into shared memory (lane -> 16 bytes); gn_wait before reading them
__device__ __forceinline__ void gn_prefetch(float* gn_s, const float* __restrict__ lnw, int warp, int lane) {
  __syncwarp();       // the warp's reads of the previous gains are done
  const int k = warp * (16 * VPL) + lane * (VPL / 2);
  if constexpr (VPL == 8)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(gn_s + k)), "l"(lnw + k) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(gn_s + k)), "l"(lnw + k) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void gn_wait() {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
}

// The weight stream of one CTA: every iteration walks the L layers' units, iterations >= head_from_iter walk the
// head units after them.  Units are copied in stream order; the cursor lives in the control block.
struct Stream {
  const uint8_t* base; uint64_t policy;
  int per_iter_layers;      // L * units per layer
  int nhead;
  int head_from_iter;       // iterations >= this one include the head units
  int n_iters;
};
// copy the next unit of the stream into `slot` (one thread); no-op past the end of the stream
__device__ __forceinline__ void stream_issue_next(const Stream& sm, Ctl* ctl, int slot, uint8_t* ring, uint64_t* full) {
#if defined(T3_ABL_NOUNITS) || defined(T3_ABL_NOSTREAM)
  return;
#endif
  const int it = ctl->next_iter;
  if (it >= sm.n_iters) return;
  const int idx = ctl->next_idx;
  mbar_arrive_expect_tx(&full[slot], T3_UNIT_BYTES);
  tma_bulk_g2s(ring + slot * T3_UNIT_BYTES, sm.base + (size_t)idx * T3_UNIT_BYTES, T3_UNIT_BYTES, &full[slot], sm.policy);
  const int end = sm.per_iter_layers + (it >= sm.head_from_iter ? sm.nhead : 0);
#if T3_PF_DIST > 0
  // The ring (128 KB) holds about half a layer of this CTA's weights, and all CTAs re-arm in the same microsecond
  // (the hand-overs keep them in lockstep), so HBM sees bursts it cannot serve in time.  The unit T3_PF_DIST ahead is
  // therefore pulled into L2 now: the ring's own copies then hit L2, and HBM streams a layer or two ahead of the
  // dependency chain.  (Next iteration = same addresses; a wrong guess at the prefill/decode seam is harmless.)
  {
    int pf = idx + T3_PF_DIST;
    if (pf >= end) pf -= end;
#ifdef T3_PF_HINT
    if (pf < end) l2_prefetch_bulk(sm.base + (size_t)pf * T3_UNIT_BYTES, T3_UNIT_BYTES, sm.policy);
#else
    if (pf < end) l2_prefetch_bulk(sm.base + (size_t)pf * T3_UNIT_BYTES, T3_UNIT_BYTES);
#endif
  }
#endif
  if (idx + 1 == end) { ctl->next_idx = 0; ctl->next_iter = it + 1; }
  else ctl->next_idx = idx + 1;
}

// ---- consume n weight units from the ring: red[j][warp][32] = this warp's k-slice partial of unit j,
// index = row * 16 + weight row of the unit
struct RingPos { int slot; uint32_t parity; long long k; };

__device__ __forceinline__ void consume_units(int n, RingPos& rp, const uint32_t (&Bf)[VPL][2], float* red, uint8_t* ring,
                                              uint64_t* full, const Stream& sm, int warp, int lane, Ctl* ctl, T3State* st) {
  const int g = lane >> 2, tig = lane & 3;
#ifdef T3_ABL_NOUNITS
  for (int j = 0; j < n; ++j)
    if (tig < 2) { float* dst = red + (j * NCONS_WARPS + warp) * 32 + tig * 16 + g; dst[0] = 0.f; dst[8] = 0.f; }
  return;
#endif
#pragma unroll 1
  for (int j = 0; j < n; ++j) {
#if !defined(T3_ABL_NOSTREAM) && !defined(T3_ABL_NOWAIT)    // timing-only builds: MMAs on whatever the ring holds
    mbar_wait(&full[rp.slot], rp.parity, ctl, st);
#endif
    const uint4* wp = reinterpret_cast<const uint4*>(ring + rp.slot * T3_UNIT_BYTES) + warp * (VPL * 32) + lane;
    uint4 w[VPL];
#pragma unroll
    for (int s = 0; s < VPL; ++s) w[s] = wp[s * 32];
    float c[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { c[q][0] = 0.f; c[q][1] = 0.f; c[q][2] = 0.f; c[q][3] = 0.f; }
#pragma unroll
    for (int s = 0; s < VPL; ++s) mma_bf16_16816(c[s & 3], w[s].x, w[s].y, w[s].z, w[s].w, Bf[s][0], Bf[s][1]);
    if (tig < 2) {   // columns 2 tig, 2 tig + 1 = hi and lo of activation row tig
      float* dst = red + (j * NCONS_WARPS + warp) * 32 + tig * 16 + g;
      dst[0] = ((c[0][0] + c[0][1]) + (c[1][0] + c[1][1])) + ((c[2][0] + c[2][1]) + (c[3][0] + c[3][1]));
      dst[8] = ((c[0][2] + c[0][3]) + (c[1][2] + c[1][3])) + ((c[2][2] + c[2][3]) + (c[3][2] + c[3][3]));
    }
    if (++rp.slot == NSLOT) { rp.slot = 0; rp.parity ^= 1u; }
    ++rp.k;
  }
}

#ifdef T3_DIRECT
// ---- experimental: weight units straight from L2 into registers (no shared-memory ring) --------------------------
// The unit's fragments are loaded with coalesced 128-bit loads BEFORE the wait for the step's inputs (weights do not
// depend on activations), so on the critical path a unit is 8 MMAs; HBM -> L2 runs ahead by bulk L2 prefetches.
__device__ __forceinline__ uint4 ldg_w(const uint4* p, uint64_t policy) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(policy));
  return v;
}
template <int J>
__device__ __forceinline__ void direct_load(uint4 (&w)[T3_MAXU][VPL], int n, const uint8_t* ub, int lofs, uint64_t policy) {
  if (J < n) {
    const uint4* wp = reinterpret_cast<const uint4*>(ub + (size_t)J * T3_UNIT_BYTES) + lofs;
#pragma unroll
    for (int s = 0; s < VPL; ++s) w[J][s] = ldg_w(wp + s * 32, policy);
  }
}
template <int J>
__device__ __forceinline__ void direct_mma(const uint4 (&w)[T3_MAXU][VPL], int n, const uint32_t (&Bf)[VPL][2], float* red,
                                           int warp, int g, int tig) {
  if (J < n) {
    float c[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { c[q][0] = 0.f; c[q][1] = 0.f; c[q][2] = 0.f; c[q][3] = 0.f; }
#pragma unroll
    for (int s = 0; s < VPL; ++s) mma_bf16_16816(c[s & 3], w[J][s].x, w[J][s].y, w[J][s].z, w[J][s].w, Bf[s][0], Bf[s][1]);
    if (tig < 2) {
      float* dst = red + (J * NCONS_WARPS + warp) * 32 + tig * 16 + g;
      dst[0] = ((c[0][0] + c[0][1]) + (c[1][0] + c[1][1])) + ((c[2][0] + c[2][1]) + (c[3][0] + c[3][1]));
      dst[8] = ((c[0][2] + c[0][3]) + (c[1][2] + c[1][3])) + ((c[2][2] + c[2][3]) + (c[3][2] + c[3][3]));
    }
  }
}
#endif

__device__ __forceinline__ float red_sum(const float* red, int j, int o) {
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < NCONS_WARPS; ++w) s += red[(j * NCONS_WARPS + w) * 32 + o];
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
  static __device__ __forceinline__ void lds8(const __nv_bfloat16* p, float (&f)[8]) {     // same from shared memory
    uint4 v = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
};
template <> struct KV<float> {
  static __device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
    float4 a = __ldcg(reinterpret_cast<const float4*>(p)), b = __ldcg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
  static __device__ __forceinline__ float round(float x) { return x; }
  static __device__ __forceinline__ void store(float* p, float x) { *p = x; }
  static __device__ __forceinline__ void lds8(const float* p, float (&f)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
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

__device__ __forceinline__ int attn_splits(int G, int B) {
  int S = G / (B * T3_NH);
  return S > MAXS ? MAXS : S;
}

// ---- phase B: RoPE + KV append + split-KV attention over the slab -------------------------------
// (reference: HF LlamaAttention eager path: softmax(q k^T / 8 + causal) v in fp32; rotate_half RoPE)
// CTA c < S*npairs handles pair = c % npairs (CFG row b, head) and KV split c / npairs; it publishes
// {sum_i exp(s_i - m) v_i, m, l} through the LL part buffer.  The S partials of a pair are merged by
// their consumers (the o_proj prologue), so there is no separate combine hop.
#ifdef T3_TIMELINE
#define T3_ATL(k) do { if (p.prof && tid == 0 && (p.prof_it & 0x40000000) && it == (p.prof_it & 0xffff)) \
                         p.prof[1024 + (size_t)c * 1024 + 512 + 4 * layer + (k)] = gtime_ns(); } while (0)
#else
#define T3_ATL(k) do {} while (0)
#endif
template <typename KVT>
__device__ __noinline__ void attn_phase_stream(const T3Params& p, int it, int layer, int pos0, float* scr, int warp, int lane,
                                        Ctl* ctl) {
  const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x, L = p.n_layers;
  const int npairs = p.B * T3_NH;
  const int S = attn_splits(G, p.B);
  if (c >= S * npairs) return;
#ifdef T3_ABL_NOATTN
  return;
#endif
  const int pair = c % npairs, split = c / npairs, b = pair / T3_NH, head = pair % T3_NH;
  float* q_s = scr;                       // [64] (pre-scaled by 1/8)
  float* k_new = scr + 64;                // [64]
  float* v_new = scr + 128;               // [64]
  float* wpart = scr + 192;               // [warps][PART_STRIDE]
  KVT* kc = reinterpret_cast<KVT*>(p.kcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;
  KVT* vc = reinterpret_cast<KVT*>(p.vcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;

  // The keys / values of the past positions do not depend on this step: their copies into shared memory
  // (cp.async, no registers: the caller's live state leaves this function few) are issued BEFORE the wait for
  // q, k, v, so their L2 / HBM latency hides behind the qkv hand-over instead of following it.  Every thread
  // copies exactly the 16-byte pieces it reads back itself: no barrier, only its own cp.async.wait_all.
  const int lg = lane >> 3, sub = lane & 7;
  const int chunk = (pos0 + S - 1) / S;
  const int lo = split * chunk;
  const int hi = min(pos0, lo + chunk);
  constexpr int PIECES = sizeof(KVT) * 8 / 16;            // 16-byte pieces per 8 elements (1 bf16, 2 fp32)
  constexpr int APF = (sizeof(KVT) == 2) ? 3 : 1;         // prefetched rounds of 4 positions per warp
  static_assert((192 + NCONS_WARPS * PART_STRIDE) * 4 <= 4096 && 4096 + APF * NCONS * 2 * PIECES * 16 <= SCRATCH_BYTES, "attention scratch");
  uint8_t* pf = reinterpret_cast<uint8_t*>(scr) + 4096 + (size_t)tid * (2 * PIECES * 16);   // [round][thread][K|V][pieces]
  // round r = positions lo + 32 r + (warp * 4 + lg); a ring of APF rounds stays in flight: round r + APF is issued
  // into the slot round r just left, one commit group per round (empty groups keep the count uniform)
  auto issue_round = [&](int r) {
    const int i = lo + warp * 4 + lg + r * (NCONS_WARPS * 4);
    if (i < hi) {
      const uint32_t dst = smem_u32(pf + (size_t)(r % APF) * (NCONS * 2 * PIECES * 16));
      const KVT* ks = kc + (size_t)i * 64 + sub * 8;
      const KVT* vs = vc + (size_t)i * 64 + sub * 8;
#pragma unroll
      for (int q = 0; q < PIECES; ++q) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + q * 16), "l"(reinterpret_cast<const uint8_t*>(ks) + q * 16) : "memory");
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (PIECES + q) * 16), "l"(reinterpret_cast<const uint8_t*>(vs) + q * 16) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int rounds = (max(hi - lo, 0) + NCONS_WARPS * 4 - 1) / (NCONS_WARPS * 4);
#pragma unroll
  for (int n = 0; n < APF; ++n) issue_round(n);
  float kr = 0.f, vr = 0.f;
  if (tid < 64) {
    const int d = tid;
    const float cs = __ldg(p.rope_cos + (size_t)pos0 * 32 + (d & 31));
    const float sn = __ldg(p.rope_sin + (size_t)pos0 * 32 + (d & 31));
    const unsigned long long* row = p.ll + LL_QKV + (size_t)b * 3072 + head * 64;
    const unsigned long long* ptr[5] = {row + d, row + (d ^ 32), row + 1024 + d, row + 1024 + (d ^ 32),
                                              row + 2048 + d};
    unsigned long long w[5];
    ll_poll<5>(ptr, 5, ll_flag(it, L, layer, LLP_QKV), w, ctl, p.st);
    const float q = ll_val(w[0]), qo = ll_val(w[1]), k = ll_val(w[2]), ko = ll_val(w[3]), v = ll_val(w[4]);
    const float qr = q * cs + (d < 32 ? -qo : qo) * sn;
    kr = KV<KVT>::round(k * cs + (d < 32 ? -ko : ko) * sn);
    vr = KV<KVT>::round(v);
    q_s[d] = qr * 0.125f;
    k_new[d] = kr;
    v_new[d] = vr;
  }
  T3_ATL(0);
  cbar();
  T3_ATL(1);

  float qv[8];
  Osm st;
  osm_init(st);
#pragma unroll
  for (int e = 0; e < 8; ++e) qv[e] = q_s[sub * 8 + e];
#pragma unroll 1
  for (int r = 0; r < rounds; ++r) {
    asm volatile("cp.async.wait_group %0;" ::"n"(APF - 1) : "memory");      // round r has landed (groups complete in order)
    const int i = lo + warp * 4 + lg + r * (NCONS_WARPS * 4);
    const bool valid = i < hi;
    float kf[8], vf[8];
    if (valid) {
      const KVT* src = reinterpret_cast<const KVT*>(pf + (size_t)(r % APF) * (NCONS * 2 * PIECES * 16));
      KV<KVT>::lds8(src, kf);
      KV<KVT>::lds8(src + 8, vf);
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) { kf[e] = 0.f; vf[e] = 0.f; }
    }
    issue_round(r + APF);          // the slot is free: this thread has its values in registers
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) s += qv[e] * kf[e];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (valid) osm_update(st, s, vf);
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  // the key of this position itself: last split, warp 0, lane group 0
  if (split == S - 1 && warp == 0) {
    float kf[8], vf[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { kf[e] = k_new[sub * 8 + e]; vf[e] = v_new[sub * 8 + e]; }
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) s += qv[e] * kf[e];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (lg == 0) osm_update(st, s, vf);
  }
  // merge the 4 lane groups, then the warps
  osm_merge_xor(st, 8);
  osm_merge_xor(st, 16);
  if (lg == 0) {
    float* w = wpart + warp * PART_STRIDE;
#pragma unroll
    for (int e = 0; e < 8; ++e) w[sub * 8 + e] = st.o[e];
    if (sub == 0) { w[64] = st.m; w[65] = st.l; }
  }
  T3_ATL(2);
  cbar();
  if (tid < 64) {
    const int d = tid;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NCONS_WARPS; ++w) M = fmaxf(M, wpart[w * PART_STRIDE + 64]);
    float Ls = 0.f, O = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < NCONS_WARPS; ++w) {
        const float* wp = wpart + w * PART_STRIDE;
        const float cw = (wp[64] == -INFINITY) ? 0.f : __expf(wp[64] - M);
        Ls += cw * wp[65];
        O += cw * wp[d];
      }
    }
    const unsigned fl = ll_flag(it, L, layer, LLP_PART);
    unsigned long long* out = p.ll + LL_PART + ((size_t)pair * MAXS + split) * PART_STRIDE;
    ll_store(out + d, O, fl);
    if (d == 0) { ll_store(out + 64, M, fl); ll_store(out + 65, Ls, fl); }
    // KV append after the partial is on its way; the fence orders it before this CTA's later LL stores
    if (split == S - 1) {
      KV<KVT>::store(kc + (size_t)pos0 * 64 + d, kr);
      KV<KVT>::store(vc + (size_t)pos0 * 64 + d, vr);
      __threadfence();
    }
  }
}

// ---- phase B, short contexts (bf16 slab, <= ATT_TP past positions per split): two-pass softmax on a tile ----
// Same contract as attn_phase_stream.  The split's K/V rows are copied into shared memory before the wait for q (as
// above) but as plain [position][64] tiles, the position's own key/value becomes one more row, and the work is laid
// out to shorten the dependency chain instead of streaming: (1) every score at once, 4 threads per position;
// (2) each warp takes the max / exponentials / sum of ALL positions redundantly (same order -> bit-identical in
// all warps, so no barrier is needed to agree on m and l); (3) warp w accumulates output dims 8w..8w+7 over all
// positions (a lane owns positions lane + 32 k and their p in registers) and a 9-shuffle reduce-scatter leaves
// one dim in every fourth lane, which publishes it.  Two CTA barriers, no online rescaling, no cross-warp merge.
constexpr int ATT_RS = 72;                       // tile row stride in bf16 (144 B: 16-byte column reads of 32 rows are conflict-free)
constexpr int ATT_SCAP = (ATT_TP + 1 + 31) / 32 * 32;          // score slots
constexpr int ATT_TILE_OFF = (64 + ATT_SCAP) * 4;              // bytes: q_s [64] fp32, scores [ATT_SCAP] fp32 before the tiles
static_assert(ATT_TILE_OFF % 16 == 0 && ATT_TILE_OFF + 2 * (ATT_TP + 1) * ATT_RS * 2 <= SCRATCH_BYTES, "attention tiles fit the scratch area");
__device__ __noinline__ void attn_phase_tile(const T3Params& p, int it, int layer, int pos0, float* scr, int warp, int lane,
                                             Ctl* ctl) {
  const int G = gridDim.x, c = blockIdx.x, tid = threadIdx.x, L = p.n_layers;
  const int npairs = p.B * T3_NH;
  const int S = attn_splits(G, p.B);
  if (c >= S * npairs) return;
#ifdef T3_ABL_NOATTN
  return;
#endif
  const int pair = c % npairs, split = c / npairs, b = pair / T3_NH, head = pair % T3_NH;
  float* q_s = scr;                        // [64], pre-scaled by 1/8
  float* s_sm = scr + 64;                  // [128] scores
  __nv_bfloat16* kt = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<uint8_t*>(scr) + ATT_TILE_OFF);
  __nv_bfloat16* vt = kt + (ATT_TP + 1) * ATT_RS;
  __nv_bfloat16* kc = reinterpret_cast<__nv_bfloat16*>(p.kcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;
  __nv_bfloat16* vc = reinterpret_cast<__nv_bfloat16*>(p.vcache) + ((size_t)(layer * 2 + b) * T3_NH + head) * p.max_pos * T3_HD;
  const int chunk = (pos0 + S - 1) / S;
  const int lo = split * chunk;
  const int hi = min(pos0, lo + chunk);
  const int n = max(hi - lo, 0);           // past positions of this split (<= ATT_TP, checked by the caller)
  const bool last = split == S - 1;
  const int nrows = n + (last ? 1 : 0);
  // 1. tiles on their way (16-byte pieces: row = piece / 8)
  for (int ch = tid; ch < n * 8; ch += NCONS) {
    const int r = ch >> 3, c8 = ch & 7;
    const uint32_t dk = smem_u32(kt + r * ATT_RS + c8 * 8), dv = smem_u32(vt + r * ATT_RS + c8 * 8);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dk), "l"(kc + (size_t)(lo + r) * 64 + c8 * 8) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dv), "l"(vc + (size_t)(lo + r) * 64 + c8 * 8) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  // 2. q, k, v of this position: RoPE; the new key / value row joins the tile of the last split
  float kr = 0.f, vr = 0.f;
  if (tid < 64) {
    const int d = tid;
    const float cs = __ldg(p.rope_cos + (size_t)pos0 * 32 + (d & 31));
    const float sn = __ldg(p.rope_sin + (size_t)pos0 * 32 + (d & 31));
    const unsigned long long* row = p.ll + LL_QKV + (size_t)b * 3072 + head * 64;
    const unsigned long long* ptr[5] = {row + d, row + (d ^ 32), row + 1024 + d, row + 1024 + (d ^ 32),
                                              row + 2048 + d};
    unsigned long long w[5];
    ll_poll<5>(ptr, 5, ll_flag(it, L, layer, LLP_QKV), w, ctl, p.st);
    const float q = ll_val(w[0]), qo = ll_val(w[1]), k = ll_val(w[2]), ko = ll_val(w[3]), v = ll_val(w[4]);
    const float qr = q * cs + (d < 32 ? -qo : qo) * sn;
    kr = KV<__nv_bfloat16>::round(k * cs + (d < 32 ? -ko : ko) * sn);
    vr = KV<__nv_bfloat16>::round(v);
    q_s[d] = qr * 0.125f;
    if (last) { kt[n * ATT_RS + d] = __float2bfloat16_rn(kr); vt[n * ATT_RS + d] = __float2bfloat16_rn(vr); }
  }
  T3_ATL(0);
  asm volatile("cp.async.wait_all;" ::: "memory");
  cbar();
  T3_ATL(1);
  // 3. scores: thread -> (position tid >> 2 (+ 64), 16 dims)
  {
    const int qd = tid & 3;
    float qv[16];
#pragma unroll
    for (int e = 0; e < 16; e += 4) {
      const float4 t4 = *reinterpret_cast<const float4*>(q_s + qd * 16 + e);
      qv[e] = t4.x; qv[e + 1] = t4.y; qv[e + 2] = t4.z; qv[e + 3] = t4.w;
    }
_Pragma(T3_STR(unroll T3_ATT_UNROLL))
    for (int i0 = 0; i0 < nrows; i0 += 64) {
      const int i = (tid >> 2) + i0;
      float acc = 0.f;
      if (i < nrows) {
        float kf[8];
        KV<__nv_bfloat16>::lds8(kt + i * ATT_RS + qd * 16, kf);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc += qv[e] * kf[e];
        KV<__nv_bfloat16>::lds8(kt + i * ATT_RS + qd * 16 + 8, kf);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc += qv[8 + e] * kf[e];
      }
      acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      acc += __shfl_xor_sync(0xffffffffu, acc, 2);
      if (qd == 0 && i < nrows) s_sm[i] = acc;
    }
  }
  T3_ATL(2);
  cbar();
  // 4. max, then exponentials / their sum / output dims 8 warp .. 8 warp + 7 in one sweep: every warp over all rows
  //    (lane owns rows lane + 32 k)
  float M = -INFINITY;
#pragma unroll 4
  for (int k0 = 0; k0 < nrows; k0 += 32) {
    const int i = lane + k0;
    M = fmaxf(M, (i < nrows) ? s_sm[i] : -INFINITY);
  }
  M = warp_max(M);
  // alignment analyzer: the conditional row of the analysed layer publishes exp(s - m) of the text columns of this split
  // and the split's (m, l); CTA 0 normalises across splits and averages the heads before it samples the next token
  const bool al_pub = layer == p.al_layer && b == 0 && warp == 0;
  const unsigned al_fl = ll_flag(it, L, layer, LLP_ALIGN);
  float lsum = 0.f;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
_Pragma(T3_STR(unroll T3_ATT_UNROLL))
  for (int k0 = 0; k0 < nrows; k0 += 32) {
    const int i = lane + k0;
    if (i < nrows) {
      const float pk = __expf(s_sm[i] - M);
      lsum += pk;
      float vf[8];
      KV<__nv_bfloat16>::lds8(vt + i * ATT_RS + warp * 8, vf);
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += pk * vf[e];
    }
  }
  lsum = warp_sum(lsum);
  if (al_pub) {        // one warp of 16 CTAs, one layer per token: kept out of the accumulation loop above
    const int c0 = max(p.al_i - lo, 0), c1 = min(p.al_i + p.al_n - lo, n);      // rows of this split that are text columns
    for (int i = c0 + lane; i < c1; i += 32)
      ll_store(p.ll + LL_AL + (size_t)head * AL_CAP + (lo + i - p.al_i), __expf(s_sm[i] - M), al_fl);
    if (lane == 0) {
      ll_store(p.ll + LL_ALML + (size_t)(head * MAXS + split) * 2, M, al_fl);
      ll_store(p.ll + LL_ALML + (size_t)(head * MAXS + split) * 2 + 1, lsum, al_fl);
    }
  }
  // reduce-scatter over the 32 lanes: 8 -> 4 -> 2 -> 1 values, then two plain steps; lane 4 j ends with dim j
  {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
    float a4[4], a2[2], a1;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float send = h16 ? acc[e] : acc[4 + e];
      const float got = __shfl_xor_sync(0xffffffffu, send, 16);
      a4[e] = (h16 ? acc[4 + e] : acc[e]) + got;
    }
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float send = h8 ? a4[e] : a4[2 + e];
      const float got = __shfl_xor_sync(0xffffffffu, send, 8);
      a2[e] = (h8 ? a4[2 + e] : a4[e]) + got;
    }
    {
      const float send = h4 ? a2[0] : a2[1];
      const float got = __shfl_xor_sync(0xffffffffu, send, 4);
      a1 = (h4 ? a2[1] : a2[0]) + got;
    }
    a1 += __shfl_xor_sync(0xffffffffu, a1, 2);
    a1 += __shfl_xor_sync(0xffffffffu, a1, 1);
    const unsigned fl = ll_flag(it, L, layer, LLP_PART);
    unsigned long long* out = p.ll + LL_PART + ((size_t)pair * MAXS + split) * PART_STRIDE;
    if ((lane & 3) == 0) ll_store(out + warp * 8 + (lane >> 2), a1, fl);
    if (warp == 0 && lane == 1) { ll_store(out + 64, M, fl); ll_store(out + 65, lsum, fl); }
  }
  // KV append after the partial is on its way; the fence orders it before this CTA's later LL stores
  if (last && tid < 64) {
    kc[(size_t)pos0 * 64 + tid] = __float2bfloat16_rn(kr);
    vc[(size_t)pos0 * 64 + tid] = __float2bfloat16_rn(vr);
    __threadfence();
  }
}

// o_proj prologue: merge the S split-KV partials of (row b, head) for this lane's k pairs.  Pair r of the lane
// sits in head (kq >> 6) + (r >> 1) at d = (kq & 63) + 32 (r & 1): the 128-wide slice of a warp spans HPS heads.
constexpr int HPS = (16 * VPL) / 64;      // heads per warp slice (2, or 1 with 16 warps)
static_assert(HPS == 1 || HPS == 2, "slice layout");
// (S == MAXS on every supported device: cbx_t3_create requires >= 129 SMs, so S = min(MAXS, G / (B 16)) = MAXS; the
// split count is a compile-time constant here and the merge is straight-line code.)
__device__ __forceinline__ void attn_combine8(float (&x)[VPL], const T3Params& p, int b, int k, unsigned flag, bool active,
                                              Ctl* ctl) {
  unsigned long long wo[MAXS][VPL], wm[MAXS][HPS], wl[MAXS][HPS];
  const int head = k >> 6, d = k & 63;       // first head of the slice; d = 2 (lane & 15)
  const unsigned long long* bm = p.ll + LL_PART + ((size_t)(b * T3_NH + head) * MAXS) * PART_STRIDE;
  const unsigned long long* bo = bm + d;
  if (!active) {
#pragma unroll
    for (int e = 0; e < VPL; ++e) x[e] = 0.f;
    return;
  }
  unsigned spins = 0; unsigned long long t0 = 0;
  for (;;) {
    bool ok = true;
#pragma unroll
    for (int s = 0; s < MAXS; ++s) {
      const unsigned long long* po = bo + s * PART_STRIDE;
      const unsigned long long* pm = bm + s * PART_STRIDE;
      ll_ld2_off<0>(po, wo[s][0], wo[s][1]);
      ll_ld2_off<32 * 8>(po, wo[s][2], wo[s][3]);
      ll_ld2_off<64 * 8>(pm, wm[s][0], wl[s][0]);
      if constexpr (VPL == 8) {     // the next head of the same row lies MAXS * PART_STRIDE words further
        ll_ld2_off<(MAXS * PART_STRIDE) * 8>(po, wo[s][VPL - 4], wo[s][VPL - 3]);
        ll_ld2_off<(MAXS * PART_STRIDE + 32) * 8>(po, wo[s][VPL - 2], wo[s][VPL - 1]);
        ll_ld2_off<(MAXS * PART_STRIDE + 64) * 8>(pm, wm[s][HPS - 1], wl[s][HPS - 1]);
      }
    }
#pragma unroll
    for (int s = 0; s < MAXS; ++s) {
#pragma unroll
      for (int e = 0; e < VPL; ++e) ok = ok && ll_ok(wo[s][e], flag);
#pragma unroll
      for (int h = 0; h < HPS; ++h) ok = ok && ll_ok(wm[s][h], flag) && ll_ok(wl[s][h], flag);
    }
    if (ok) break;
    if (poll_timeout(spins, t0, ctl, p.st)) break;
  }
#pragma unroll
  for (int h = 0; h < HPS; ++h) {
    float Mx = -INFINITY;
#pragma unroll
    for (int s = 0; s < MAXS; ++s) Mx = fmaxf(Mx, ll_val(wm[s][h]));
    float Lt = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < MAXS; ++s) {
      const float ms = ll_val(wm[s][h]);
      const float cw = (ms == -INFINITY) ? 0.f : __expf(ms - Mx);
      Lt += cw * ll_val(wl[s][h]);
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[e] += cw * ll_val(wo[s][(VPL == 8 ? 4 * h : 0) + e]);
    }
    const float inv = (Lt > 0.f) ? 1.0f / Lt : 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) x[(VPL == 8 ? 4 * h : 0) + e] = acc[e] * inv;
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

// ---- AlignmentStreamAnalyzer.step on CTA 0 (alignment_stream_analyzer.py:89-154) ---------------------------------
// Runs before the pending logits are processed: builds the head-mean attention row(s) of the forward pass that
// produced those logits over the text columns — from the prompt prefill (first call), from the LL words the analysed
// layer's attention CTAs published (decode), or from an injected script (tests) — applies the monotonic masking,
// updates the running statistics and leaves the decision (force / suppress EOS) in the shared-memory state.
// `row` = the sampling scratch (free at this point): [AL_CAP] row, then 64 floats of reduction scratch, then (m, l) data.
__device__ __noinline__ void align_step(const T3Params& p, int it, int pos0, float* row, AlState* as, Ctl* ctl, int warp,
                                        int lane) {
  const int tid = threadIdx.x, S = p.al_n, L = p.n_layers;
  float* rs = row + AL_CAP;
  int* ri = reinterpret_cast<int*>(rs + 32);
  float* ml = rs + 64;                     // [16][MAXS][2] split (m, l), then [16][2] = {M_h, 1 / (16 L_h)}
  const int nrows = (as->frame == 0) ? p.al_rows0 : 1;
  for (int r = 0; r < nrows; ++r) {
    const int T = as->T;                   // rows consumed so far (uniform: written by thread 0 behind a barrier)
    if (p.al_inject) {
      for (int c = tid; c < S; c += NCONS) row[c] = __ldg(p.al_inject + (size_t)T * S + c);
    } else if (as->frame == 0) {
      for (int c = tid; c < S; c += NCONS) row[c] = __ldcg(p.al_prefill + (size_t)r * AL_CAP + c);
    } else {
      const unsigned fl = ll_flag(it - 1, L, p.al_layer, LLP_ALIGN);
      if (tid < T3_NH * MAXS * 2) {
        const unsigned long long* wp = p.ll + LL_ALML + tid;
        unsigned long long w;
        unsigned spins = 0; unsigned long long t0 = 0;
        for (;;) { w = ll_ld1(wp); if (ll_ok(w, fl)) break; if (poll_timeout(spins, t0, ctl, p.st)) break; }
        ml[tid] = ll_val(w);
      }
      cbar();
      if (tid < T3_NH) {
        float Mt = -INFINITY;
#pragma unroll
        for (int s = 0; s < MAXS; ++s) Mt = fmaxf(Mt, ml[(tid * MAXS + s) * 2]);
        float Lt = 0.f;
#pragma unroll
        for (int s = 0; s < MAXS; ++s) {
          const float ms = ml[(tid * MAXS + s) * 2];
          Lt += (ms == -INFINITY) ? 0.f : ml[(tid * MAXS + s) * 2 + 1] * __expf(ms - Mt);
        }
        ml[T3_NH * MAXS * 2 + tid * 2] = Mt;
        ml[T3_NH * MAXS * 2 + tid * 2 + 1] = 1.0f / (Lt * (float)T3_NH);
      }
      cbar();
      const int chunk = (pos0 - 1 + MAXS - 1) / MAXS;      // KV split size of the forward pass that produced the logits
      for (int c = tid; c < S; c += NCONS) {
        const int s = min((p.al_i + c) / chunk, MAXS - 1);
        float acc = 0.f;
        for (int h = 0; h < T3_NH; ++h) {
          const unsigned long long* wp = p.ll + LL_AL + (size_t)h * AL_CAP + c;
          unsigned long long w;
          unsigned spins = 0; unsigned long long t0 = 0;
          for (;;) { w = ll_ld1(wp); if (ll_ok(w, fl)) break; if (poll_timeout(spins, t0, ctl, p.st)) break; }
          acc += ll_val(w) * __expf(ml[(h * MAXS + s) * 2] - ml[T3_NH * MAXS * 2 + h * 2]) * ml[T3_NH * MAXS * 2 + h * 2 + 1];
        }
        row[c] = acc;
      }
    }
    cbar();
    if (p.al_rows_out)
      for (int c = tid; c < S; c += NCONS) p.al_rows_out[(size_t)T * S + c] = row[c];
    // monotonic masking (:107), arg-max (first index on ties), max over columns [:S-5]
    const int frame = as->frame;
    float amax = -INFINITY, rmax = -INFINITY; int aidx = 0x7fffffff;
    for (int c = tid; c < S; c += NCONS) {
      const float v = (c >= frame + 1) ? 0.f : row[c];
      row[c] = v;
      if (v > amax) { amax = v; aidx = c; }
      if (c < S - 5) rmax = fmaxf(rmax, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, amax, o);
      const int i2 = __shfl_xor_sync(0xffffffffu, aidx, o);
      if (m2 > amax || (m2 == amax && i2 < aidx)) { amax = m2; aidx = i2; }
      rmax = fmaxf(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
    }
    cbar();
    if (lane == 0) { rs[warp] = amax; ri[warp] = aidx; rs[NCONS_WARPS + warp] = rmax; }
    cbar();
    if (tid == 0) {
      float A = rs[0], R = rs[NCONS_WARPS]; int I = ri[0];
      for (int w = 1; w < NCONS_WARPS; ++w) {
        if (rs[w] > A || (rs[w] == A && ri[w] < I)) { A = rs[w]; I = ri[w]; }
        R = fmaxf(R, rs[NCONS_WARPS + w]);
      }
      as->T = T + 1;
      as->last2_prev = as->last2_cur;
      as->last2_cur = fmaxf(row[S - 2], row[S - 1]);
      float f4 = as->first4;
      for (int c = 0; c < 4 && c < S; ++c) f4 = fmaxf(f4, row[c]);
      as->first4 = f4;
      if (as->complete) {                    // rows at index >= completed_at
        as->tail0 += row[S - 3]; as->tail1 += row[S - 2]; as->tail2 += row[S - 1];
        if (S > 5) as->rep += R;
      }
      if (r == nrows - 1) {
        const int cur = I, d = cur - as->text_pos;
        if (d > -4 && d < 7) as->text_pos = cur;                                   // not a discontinuity (:118-121)
        const float a2 = fmaxf(as->last2_prev, as->last2_cur);
        const bool false_start = !as->started && (a2 > 0.1f || as->first4 < 0.5f);   // (:126)
        as->started = false_start ? 0 : 1;
        if (as->text_pos >= S - 3) as->complete = 1;                               // (:131)
        const bool long_tail = as->complete && fmaxf(as->tail0, fmaxf(as->tail1, as->tail2)) >= 10.f;   // (:140)
        const bool repetition = as->complete && as->rep > 5.f;                     // (:143)
        as->force = (long_tail || repetition) ? 1 : 0;
        as->suppress = (cur < S - 3) ? 1 : 0;                                      // (:153)
        as->frame = frame + 1;
        if (p.al_out) {
          p.al_out[it * 4 + 0] = as->text_pos;
          p.al_out[it * 4 + 1] = as->started | (as->complete << 1) | (as->force << 2) | (as->suppress << 3);
          p.al_out[it * 4 + 2] = cur;
          p.al_out[it * 4 + 3] = as->T;
        }
      }
    }
    cbar();
  }
}

// Processes the pending logits of the CFG pair and emits token `it`.  All NCONS threads of CTA 0.
#ifdef T3_PROFILE
#define T3_SPROF(k) do { if (p.prof && threadIdx.x == 0 && it == p.prof_it) p.prof[300 + (k)] = gtime_ns(); } while (0)
#else
#define T3_SPROF(k) do {} while (0)
#endif
__device__ __noinline__ void sample_phase(const T3Params& p, int it, int n_gen, float* sl, unsigned* seen, Ctl* ctl, int warp,
                             int lane, int pos0, AlState* as) {
  const int tid = threadIdx.x, L = p.n_layers;
  bool al_force = false, al_sup = false;
  if (p.al_layer >= 0) {
    align_step(p, it, pos0, sl, as, ctl, warp, lane);
    al_force = as->force != 0; al_sup = as->suppress != 0;
    cbar();          // every thread has its copy of the decision before sl[] is reused for the logits
  }
  const bool al_any = al_force || al_sup;
  float* rs = sl + VOCAB_PAD;          // reduction scratch (64 floats)
  int* ri = reinterpret_cast<int*>(rs + 32);
  const CbxSampleParams& sp = p.sp;
  const bool cfg = sp.cfg_weight > 0.f && p.B == 2;
  T3_SPROF(0);
  // 1. CFG, temperature, repetition penalty; running max (first index on ties)
  float lmax = -INFINITY; int imax = 0x7fffffff;
#ifdef T3_ABL_NOSAMPLE
  constexpr int NK = 0;
#else
  constexpr int NK = (T3_VOCAB + NCONS - 1) / NCONS;   // 33
#endif
  const unsigned lfl = ll_flag(it - 1, L, L, LLP_LOGITS);
  // Every field this loop needs is copied into registers first: `p` is a shared-memory copy, and after each polling
  // asm (a compiler barrier) or store to sl[] the compiler would have to reload p.B, p.ll, sp.* from it again, each a
  // dependent shared-memory round trip in front of the next load (measured: 16 us of the 18 us of this phase).
  const int B2 = p.B == 2;
  const unsigned long long* llc = p.ll + LL_LOGITS;
  const unsigned long long* llu = llc + VOCAB_PAD;
  const float* plog = p.logits;
  float* lout = p.logits_out;
  const float cfgw = sp.cfg_weight, temp = sp.temperature, rpen = sp.repetition_penalty;
  T3State* stp = p.st;
  constexpr int KB = T3_SAMPLE_KB;       // words per row in flight per thread (33 = 3 batches of 11)
  for (int k0 = 0; k0 < NK; k0 += KB) {
    float cv[KB], uv[KB];
    if (it == 0) {   // logits of the previous launch (prefill or an earlier decode call)
#pragma unroll
      for (int k = 0; k < KB; ++k) {
        const int i = tid + (k0 + k) * NCONS;
        cv[k] = (i < T3_VOCAB) ? __ldcg(plog + i) : 0.f;
        uv[k] = (i < T3_VOCAB && B2) ? __ldcg(plog + T3_VOCAB + i) : 0.f;
      }
    } else {
      unsigned long long wc[KB], wu[KB];
      unsigned spins = 0; unsigned long long t0 = 0;
      for (;;) {
        bool ok = true;
#pragma unroll
        for (int k = 0; k < KB; ++k) {
          const int i = min(tid + (k0 + k) * NCONS, T3_VOCAB - 1);    // tail threads re-read the last word: no branch
          wc[k] = ll_ld1(llc + i);
          wu[k] = ll_ld1((B2 ? llu : llc) + i);
        }
#pragma unroll
        for (int k = 0; k < KB; ++k) ok = ok && ll_ok(wc[k], lfl) && ll_ok(wu[k], lfl);
        if (ok) break;
        if (poll_timeout(spins, t0, ctl, stp)) break;
      }
#pragma unroll
      for (int k = 0; k < KB; ++k) { cv[k] = ll_val(wc[k]); uv[k] = B2 ? ll_val(wu[k]) : 0.f; }
    }
#pragma unroll
    for (int k = 0; k < KB; ++k) {
      const int i = tid + (k0 + k) * NCONS;
      if (i < T3_VOCAB) {
        const float c = cv[k], u = uv[k];
        if (lout) {
          lout[((size_t)it * (1 + B2)) * T3_VOCAB + i] = c;
          if (B2) lout[((size_t)it * 2 + 1) * T3_VOCAB + i] = u;
        }
        float cc = c, uu = u;
        // alignment analyzer: both CFG rows get the reference's overrides (alignment_stream_analyzer.py:146-153); the
        // suppression comes second and wins, as there
        if (al_any) {
          if (al_force) cc = uu = (i == STOP_SPEECH) ? 32768.f : -32768.f;
          if (al_sup && i == STOP_SPEECH) cc = uu = -32768.f;
        }
        float x = cfg ? cc + cfgw * (cc - uu) : cc;
        if (temp != 1.0f) x = x / temp;
        if ((seen[i >> 5] >> (i & 31)) & 1u) x = (x < 0.f) ? x * rpen : x / rpen;
        sl[i] = x;
        if (x > lmax) { lmax = x; imax = i; }
      }
    }
  }
  T3_SPROF(1);
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
  T3_SPROF(2);
  if (p.forced) tok = p.forced[it];
#if defined(T3_ABL_NOPOLL) || defined(T3_ABL_NOUNITS) || defined(T3_ABL_NOSTREAM) || defined(T3_ABL_NOWAIT)
  tok = (unsigned)(tok + it) % 6000u;
#endif
  const bool stop = sp.stop_on_eos && tok == STOP_SPEECH;
  if (tid == 0) {
    p.tokens_out[it] = tok;
    seen[tok >> 5] |= 1u << (tok & 31);
    p.seen[tok] = 1;
    if (stop) p.st->stop = 1;
  }
  // 5. next-token embedding: speech_emb[tok] + speech_pos_emb[n_gen + 1]  (t3.py:351-356); both CFG rows get the
  //    same vector, published once together with the control word {token | stop bit}
  if (tid < T3_H / 4) {
    const unsigned fl = ll_flag(it, L, 0, LLP_EMB);
    const float4 e = __ldg(reinterpret_cast<const float4*>(p.speech_emb + (size_t)tok * T3_H) + tid);
    const float4 q = __ldg(reinterpret_cast<const float4*>(p.speech_pos_emb + (size_t)(n_gen + 1) * T3_H) + tid);
    unsigned long long* dst = p.ll + LL_EMB + 4 * tid;
    ll_store(dst + 0, e.x + q.x, fl); ll_store(dst + 1, e.y + q.y, fl);
    ll_store(dst + 2, e.z + q.z, fl); ll_store(dst + 3, e.w + q.w, fl);
    if (tid == 0) ll_store(p.ll + LL_EMB + T3_H, __uint_as_float((unsigned)tok | (stop ? 0x80000000u : 0u)), fl);
  }
  T3_SPROF(3);
  cbar();   // seen[] update visible to the next call
}

// ---- the megakernel -----------------------------------------------------------------------------
// One iteration = one position.  Its 4 L (+1) GEMV steps — step = 4 l + phase, phase in {qkv, o, gate/up, down},
// then the head — run through ONE rolled loop body, so the code of a step is shared by every phase and layer and
// stays resident in the instruction cache; what differs between phases is data (a few switches).
template <typename KVT>
__global__ void __launch_bounds__(NTHREADS, 1) t3_megakernel(const T3Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SMEM_BARS);
  Ctl* ctl = reinterpret_cast<Ctl*>(smem + SMEM_CTL);
  float* red_all = reinterpret_cast<float*>(smem + SMEM_RED);
  float* ssq_all = reinterpret_cast<float*>(smem + SMEM_SSQ);
  float* scr = reinterpret_cast<float*>(smem + SMEM_SCRATCH);
  unsigned* seen = reinterpret_cast<unsigned*>(smem + SMEM_SEEN);
  T3Params* ps = reinterpret_cast<T3Params*>(smem + SMEM_PARAMS);
  AlState* al_s = reinterpret_cast<AlState*>(smem + SMEM_SSQ + 128);      // second half of the ssq area (the sums use 128 B)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x, c = blockIdx.x;
  const int L = p.n_layers;
  const int n_iters = (p.mode == 0) ? p.P : p.max_new;
  Stream sm;
  sm.base = p.wstream + (size_t)c * p.stride_bytes;
  sm.policy = l2_policy_evict_first();
  sm.per_iter_layers = L * t3_plan_units_per_layer(c, G);
  sm.nhead = t3_plan_count(PH_HEAD, c, G);
  sm.head_from_iter = (p.mode == 1) ? 0 : n_iters - 1;
  sm.n_iters = n_iters;
  if (tid == 32) *ps = p;
  if (tid == 64 && blockIdx.x == 0 && p.al_layer >= 0) *al_s = *p.al_state;
  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) mbar_init(&full[s], 1);
    ctl->abort = 0; ctl->next_idx = 0; ctl->next_iter = 0;
    mbar_fence_init();
#if T3_PF_DIST > 0
    for (int u = NSLOT; u < T3_PF_DIST && u < sm.per_iter_layers; ++u) l2_prefetch_bulk(sm.base + (size_t)u * T3_UNIT_BYTES, T3_UNIT_BYTES);
#endif
#ifndef T3_DIRECT
    for (int s = 0; s < NSLOT; ++s) stream_issue_next(sm, ctl, s, ring, full);
#endif
  }
  if (tid < PH_COUNT * T3_MAXU) {
    const int ph = tid / T3_MAXU, j = tid % T3_MAXU, n = t3_plan_count(ph, c, G);
    if (j == 0) ctl->plan_n[ph] = n;
    ctl->plan_u[ph][j] = j < n ? t3_plan_unit(ph, c, G, j) : 0;
  }
  if (c == 0)      // byte set in global memory (crosses launches) -> bit set: one ballot per 32 ids
    for (int i = tid; i < SEEN_WORDS * 32; i += NTHREADS) {
      const unsigned m = __ballot_sync(0xffffffffu, i < VOCAB_PAD && p.seen[i] != 0);
      if ((tid & 31) == 0) seen[i >> 5] = m;
    }
  __syncthreads();

  const int b = lane >> 4;                        // activation row (CFG cond / uncond) of this lane
  const int kq = warp * (16 * VPL) + (lane & 15) * 2;      // its first k pair in a 1024-wide vector (the others: + 32 r)
  const bool active = b < p.B;
  uint32_t* xt = reinterpret_cast<uint32_t*>(smem + SMEM_XT) + warp * XT_WORDS;
  const int S = attn_splits(G, p.B);
  const int down_slice = ctl->plan_u[PH_DOWN][0] % T3_DOWN_KS;
  RingPos rp{0, 0u, 0ll};
  int pos0 = __ldcg(&p.st->pos);
  int n_gen = __ldcg(&p.st->n_gen);
  int n_done = 0;
  // residual stream slice (replicated in every CTA): this lane's pairs, at + 32 r floats
  float* xr = reinterpret_cast<float*>(smem + SMEM_XR) + warp * (32 * VPL) + (lane >> 4) * (16 * VPL) + (lane & 15) * 2;
  float x[VPL];
  float* gn_s = reinterpret_cast<float*>(smem + SMEM_GN);     // gains of the next RMSNorm (copied a phase ahead)
  uint32_t Bf[VPL][2];
  [[maybe_unused]] int prof_k = 0;
  int par = 0;          // parity of the running step counter: red / ssq double buffer
  gn_prefetch(gn_s, p.ln1, warp, lane);
#ifndef T3_PROFILE     // phase stamps (tools/t3_phase_times.py) only in -DT3_PROFILE builds: even dead tests cost issue slots
#define T3_PROF() do {} while (0)
#else
#define T3_PROF() do { if (p.prof && c == 0 && tid == 0 && it == p.prof_it) p.prof[prof_k++] = gtime_ns(); } while (0)
#endif
// timeline build (-DT3_TIMELINE, tools/t3_timeline.py): thread 0 of EVERY CTA stamps prologue / consume / epilogue
// end of every step at prof[1024 + c * 1024 + 4 * step + k].  Off by default: even the dead test costs ~5 %.
#ifdef T3_TIMELINE
#define T3_TL(k) do { if (p.prof && tid == 0 && (p.prof_it & 0x40000000) && it == (p.prof_it & 0xffff)) \
                        p.prof[1024 + (size_t)c * 1024 + 4 * step + (k)] = gtime_ns(); } while (0)
#else
#define T3_TL(k) do {} while (0)
#endif
#ifndef T3_PROFILE
#define T3_SUB(k) do {} while (0)
#else
#define T3_SUB(k) do { if (p.prof && c == 0 && tid == 0 && it == p.prof_it && l == 5 && (ph == PH_QKV || ph == PH_GU)) \
                         p.prof[200 + (ph == PH_GU ? 8 : 0) + (k)] = gtime_ns(); } while (0)
#endif

  for (int it = 0; it < n_iters; ++it) {
    if (p.mode == 1) {
      T3_PROF();
      if (c == 0) sample_phase(*ps, it, n_gen, scr, seen, ctl, warp, lane, pos0, al_s);
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
    }
#pragma unroll
    for (int r = 0; r < VPL / 2; ++r) *reinterpret_cast<float2*>(xr + 32 * r) = make_float2(0.f, 0.f);
    const bool with_head = (p.mode == 1) || (it == n_iters - 1);
    const int n_steps = 4 * L + (with_head ? 1 : 0);
#ifdef T3_DIRECT
    int uidx = 0;        // unit index inside this CTA's per-iteration stream
    uint4 wreg[T3_MAXU][VPL];
    [[maybe_unused]] bool preloaded = false;     // -DT3_EARLY: this step's units were requested at the end of the previous step
    // request the units of a step (phase ph_, first unit u0): all of them, except the head's last two (its prologue polls
    // four pieces: they follow the poll); HBM -> L2 runs T3_PF_DIST units ahead
    auto issue_units = [&](int ph_, int u0) {
      const int n_ = ctl->plan_n[ph_];
      const uint8_t* ub = sm.base + (size_t)u0 * T3_UNIT_BYTES;
      const int lofs = warp * (VPL * 32) + lane;
      direct_load<0>(wreg, n_, ub, lofs, sm.policy);
      direct_load<1>(wreg, n_, ub, lofs, sm.policy);
      if (ph_ != PH_HEAD) {
        direct_load<2>(wreg, n_, ub, lofs, sm.policy);
        direct_load<3>(wreg, n_, ub, lofs, sm.policy);
      }
      if (tid == NCONS - 32) {
        const int end = sm.per_iter_layers + (with_head ? sm.nhead : 0);
        for (int j = 0; j < n_; ++j) {
          int pf = u0 + j + T3_PF_DIST;
          if (pf >= end) pf -= end;
          if (pf < end) l2_prefetch_bulk(sm.base + (size_t)pf * T3_UNIT_BYTES, T3_UNIT_BYTES);
        }
      }
    };
#endif

#pragma unroll 1
    for (int step = 0; step < n_steps; ++step) {
      const int l = step >> 2;
      const int ph = (l == L) ? PH_HEAD : (step & 3);
      float* red = red_all + par * RED_FLOATS;
      float* ssq = ssq_all + par * (NCONS_WARPS * 2);
#ifdef T3_DIRECT
      const int n = ctl->plan_n[ph];
#ifdef T3_EARLY
      if (!preloaded) issue_units(ph, uidx);
#else
      issue_units(ph, uidx);
#endif
#endif
      // ---------- prologue: this lane's 8 input values ----------
      T3_SUB(0);
      if (ph == PH_O) {
        // CTAs without an o_proj unit do not need the attention output at all (x only feeds the fragments)
        attn_combine8(x, p, b, kq, ll_flag(it, L, l, LLP_PART), active && ctl->plan_n[PH_O] > 0, ctl);
      } else if (ph == PH_QKV && l == 0 && p.mode == 0) {
#pragma unroll
        for (int e = 0; e < VPL; ++e) x[e] = 0.f;
        if (active) ln_fetch(x, p.embeds + ((size_t)b * p.P + it) * T3_H, kq);
      } else {
        const unsigned long long* base; size_t stride = 0; int np = 1; unsigned fl;
        if (ph == PH_GU) { base = p.ll + LL_OPROJ + (size_t)b * T3_H + kq; fl = ll_flag(it, L, l, LLP_OPROJ); }
        else if (ph == PH_DOWN) { base = p.ll + LL_ACT + (size_t)b * T3_INTER + down_slice * 1024 + kq; fl = ll_flag(it, L, l, LLP_ACT); }
        else if (l == 0) { base = p.ll + LL_EMB + kq; fl = ll_flag(it, L, 0, LLP_EMB); }          // qkv of layer 0
        else { base = p.ll + LL_DOWN + (size_t)b * T3_H + kq; stride = 2 * T3_H; np = T3_DOWN_KS; fl = ll_flag(it, L, l - 1, LLP_DOWN); }
        ll_load8(x, base, stride, np, fl, active && (ph != PH_DOWN || ctl->plan_n[PH_DOWN] > 0), ctl, p.st);
      }
      T3_SUB(1);
      T3_TL(0);
      // ---------- residual add + RMSNorm gain (the scale is applied in the epilogue) ----------
      if (ph == PH_QKV || ph == PH_GU || ph == PH_HEAD) {
#pragma unroll
        for (int r = 0; r < VPL / 2; ++r) {
          float2 v = *reinterpret_cast<const float2*>(xr + 32 * r);
          v.x += x[2 * r]; v.y += x[2 * r + 1];
          *reinterpret_cast<float2*>(xr + 32 * r) = v;
          x[2 * r] = v.x; x[2 * r + 1] = v.y;
        }
        ssq_partial(x, ssq, warp, lane);
        gn_wait();
#pragma unroll
        for (int r = 0; r < VPL / 2; ++r) {
          const float2 g2 = *reinterpret_cast<const float2*>(gn_s + kq + 32 * r);
          x[2 * r] *= g2.x; x[2 * r + 1] *= g2.y;
        }
        const float* nxt;
        if (ph == PH_QKV) nxt = p.ln2 + (size_t)l * T3_H;
        else if (ph == PH_GU) nxt = (l == L - 1 && with_head) ? p.final_norm : p.ln1 + (size_t)((l + 1) % L) * T3_H;
        else nxt = p.ln1;
        gn_prefetch(gn_s, nxt, warp, lane);
      }
      make_bfrags(Bf, x, xt, lane);
      T3_SUB(2);
#ifdef T3_DIRECT
      {
        const int g = lane >> 2, tig = lane & 3;
        if (ph == PH_HEAD) {
          const uint8_t* ub = sm.base + (size_t)uidx * T3_UNIT_BYTES;
          const int lofs = warp * (VPL * 32) + lane;
          direct_load<2>(wreg, n, ub, lofs, sm.policy);
          direct_load<3>(wreg, n, ub, lofs, sm.policy);
        }
        direct_mma<0>(wreg, n, Bf, red, warp, g, tig);
        direct_mma<1>(wreg, n, Bf, red, warp, g, tig);
        direct_mma<2>(wreg, n, Bf, red, warp, g, tig);
        direct_mma<3>(wreg, n, Bf, red, warp, g, tig);
        uidx += n;
#ifdef T3_EARLY
        // The registers of this step's units are free: request the NEXT step's units now, so their L2 -> SM transfer
        // (up to 128 KB on a 4-unit CTA, ~1 us of ingress that the polls of the next prologue would queue behind) overlaps
        // this step's reduction / epilogue and the hand-over.  Not across the attention call (128 live registers would be
        // spilled around it): the o_proj units (one per CTA) are requested at the top of their step as before.
        preloaded = false;
        if (ph != PH_QKV && step + 1 < n_steps) {
          const int l1 = (step + 1) >> 2;
          issue_units((l1 == L) ? PH_HEAD : ((step + 1) & 3), uidx);
          preloaded = true;
        }
#endif
      }
#else
      const int n = ctl->plan_n[ph];
      consume_units(n, rp, Bf, red, ring, full, sm, warp, lane, ctl, p.st);
#endif
      T3_SUB(3);
      T3_TL(1);
      cbar();
      // every warp is past its reads of the n slots of this step: the last warp (never part of the epilogue, which
      // takes warps 0..n-1) re-arms them with the units NSLOT ahead in the stream.  No per-unit release counting: the
      // unit loop stays free of atomics, and HBM has the slack (the copies are needed a whole step later).
#ifndef T3_DIRECT
      if (tid == NCONS - 32 && n > 0 && !ctl->abort) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        int sl = rp.slot - n;
        if (sl < 0) sl += NSLOT;
        for (int j = 0; j < n; ++j) {
          stream_issue_next(sm, ctl, sl, ring, full);
          if (++sl == NSLOT) sl = 0;
        }
      }
#endif
      T3_SUB(4);
      // ---------- epilogue: cross-warp sum, scale / SwiGLU, publish ----------
#ifdef T3_ABL_NOEPI
      if (false) {
#else
      if (tid < n * 32) {
#endif
        const int j = tid >> 5, o = tid & 31, bb = o >> 4, row = o & 15;
        const int u = ctl->plan_u[ph][j];
        float v = red_sum(red, j, o);
        // gate/up: rows 0-7 of the unit are gate rows, 8-15 the matching up rows: lane o pairs with lane o + 8
        // (whole-warp shuffle: kept outside the row test, B == 1 leaves lanes 16-31 without a row)
        const float other = __shfl_down_sync(0xffffffffu, v, 8);
        if (bb < p.B) {
          if (ph == PH_QKV) {
            ll_store(p.ll + LL_QKV + (size_t)bb * 3072 + u * 16 + row, v * rms_scale(ssq, bb), ll_flag(it, L, l, LLP_QKV));
          } else if (ph == PH_O) {
            ll_store(p.ll + LL_OPROJ + (size_t)bb * T3_H + u * 16 + row, v, ll_flag(it, L, l, LLP_OPROJ));
          } else if (ph == PH_GU) {
            const float sc = rms_scale(ssq, bb);
            if (row < 8) {
              const float gate = v * sc, up = other * sc;
              ll_store(p.ll + LL_ACT + (size_t)bb * T3_INTER + u * 8 + row, gate / (1.0f + __expf(-gate)) * up,
                       ll_flag(it, L, l, LLP_ACT));
            }
          } else if (ph == PH_DOWN) {
            ll_store(p.ll + LL_DOWN + (size_t)((u % T3_DOWN_KS) * 2 + bb) * T3_H + (u / T3_DOWN_KS) * 16 + row, v,
                     ll_flag(it, L, l, LLP_DOWN));
          } else {
            const int vrow = u * 16 + row;
            if (vrow < T3_VOCAB) {
              v *= rms_scale(ssq, bb);
              p.logits[(size_t)bb * T3_VOCAB + vrow] = v;     // plain copy: read by the next launch
              if (p.mode == 1) ll_store(p.ll + LL_LOGITS + (size_t)bb * VOCAB_PAD + vrow, v, ll_flag(it, L, L, LLP_LOGITS));
            }
          }
        }
      }
      T3_SUB(5);
      T3_TL(2);
      par ^= 1;
      T3_PROF();
      // ---------- RoPE, KV append, split-KV attention between the qkv and o_proj steps ----------
      if (ph == PH_QKV) {
        bool tile_path = false;
        if constexpr (sizeof(KVT) == 2) tile_path = (pos0 + S - 1) / S <= ATT_TP;
#ifdef T3_ATTN_STREAM
        tile_path = false;
#endif
        if (tile_path) attn_phase_tile(*ps, it, l, pos0, scr, warp, lane, ctl);
        else attn_phase_stream<KVT>(*ps, it, l, pos0, scr, warp, lane, ctl);
        T3_TL(3);
        T3_PROF();
      }
    }
    pos0 += 1;
    if (ctl->abort) break;
  }
  // publish state
  if (c == 0 && tid == 0) {
    p.st->pos = pos0;
    p.st->n_gen = n_gen;
    // a device error (watchdog) reaches the host in the same word as the token count: -(n + 1)
    if (p.mode == 1 && p.n_out) *p.n_out = (ctl->abort || __ldcg(&p.st->err)) ? -(n_done + 1) : n_done;
    if (p.al_layer >= 0) *p.al_state = *al_s;
  }
  __syncthreads();
  // drain: every copy in flight must have landed before the CTA may exit (early stop on EOS / abort)
  if (tid == 0) {
    // units issued so far = consumed + those still in the ring; the cursor tells how many were issued
    long long issued = 0;
    for (int i2 = 0; i2 < ctl->next_iter && i2 < n_iters; ++i2) issued += sm.per_iter_layers + (i2 >= sm.head_from_iter ? sm.nhead : 0);
    if (ctl->next_iter < n_iters) issued += ctl->next_idx;
    for (long long kk = rp.k; kk < issued; ++kk) mbar_wait(&full[kk % NSLOT], (uint32_t)((kk / NSLOT) & 1), ctl, p.st);
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
  // alignment analyzer configuration (cbx_t3_set_alignment); layer < 0 = off
  int al_layer, al_i, al_n, al_rows0;
  const float* al_inject; float* al_rows_out; int* al_out;
  float* al_prefill; AlState* al_state;
  int host_pos;         // upper bound of the cache length after the calls issued so far (EOS may stop a launch early)
  int host_gen;         // same for the generated-token count
  long long launches;
  unsigned long long* prof; int prof_it;
  void* prefill_ws;     // workspace of the GEMM prefill (NULL without plain weights)
  // the GEMM prefill of a (B, P) seen twice is captured into a CUDA graph (it is ~240 short launches: launch-bound);
  // the caller's embeddings are staged so the pointers baked into the graph stay valid
  struct PrefillGraph { int B, P, seen; cudaGraphExec_t exec; };
  std::vector<PrefillGraph> pf_graphs;
  float* pf_embeds;
  cudaStream_t cap_stream;
};

extern "C" int cbx_t3_grid(int device) {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    cbx_set_error("cbx_t3_grid: no CUDA device %d (this library has no CPU fallback)", device);
    return -1;
  }
  return prop.multiProcessorCount;
}
extern "C" int cbx_t3_kwarps(void) { return T3_KW; }
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
  CBX_ON_DEVICE(cfg->device);
  cudaDeviceProp prop;
  CBX_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  CBX_CHECK(prop.major == 10, "cbx_t3_create: built for sm_100a only, device is sm_%d%d", prop.major, prop.minor);
  const int G = prop.multiProcessorCount;
  CBX_CHECK(G >= 129 && (513 + G - 1) / G <= T3_MAXU, "cbx_t3_create: %d SMs unsupported by the plan", G);
  CBX_CHECK(G / (2 * T3_NH) >= MAXS, "cbx_t3_create: %d SMs give fewer than %d KV splits", G, MAXS);
  int coop = 0;
  CBX_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, cfg->device));
  CBX_CHECK(coop, "cbx_t3_create: device lacks cooperative launch");

  cbx_t3* h = new cbx_t3();
  h->cfg = *cfg; h->w = *w; h->G = G; h->B = 2; h->host_pos = 0; h->host_gen = 0; h->launches = 0;
  h->al_layer = -1; h->al_i = 0; h->al_n = 0; h->al_rows0 = 0; h->al_inject = nullptr; h->al_rows_out = nullptr; h->al_out = nullptr; h->prof = nullptr; h->prof_it = -1;
  const size_t kvel = cfg->kv_fp32 ? 4 : 2;
  const size_t kv_bytes = (size_t)cfg->n_layers * 2 * T3_NH * cfg->max_pos * T3_HD * kvel;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t o_k = take(kv_bytes), o_v = take(kv_bytes);
  const size_t o_ll = take(LL_WORDS * 8);
  const size_t o_logits = take(sizeof(float) * 2 * T3_VOCAB);
  const size_t o_st = take(256), o_seen = take(VOCAB_PAD);
  const size_t o_alp = take((size_t)2 * AL_CAP * sizeof(float)), o_als = take(sizeof(AlState));
  const size_t o_ws = w->wplain ? take(cbx::t3_prefill_workspace_bytes(cfg->max_pos)) : 0;
  const size_t o_pfe = w->wplain ? take((size_t)2 * cfg->max_pos * T3_H * sizeof(float)) : 0;
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
  p.al_layer = -1;
  h->al_prefill = reinterpret_cast<float*>(s + o_alp);
  h->al_state = reinterpret_cast<AlState*>(s + o_als);
  h->prefill_ws = w->wplain ? s + o_ws : nullptr;
  h->pf_embeds = w->wplain ? reinterpret_cast<float*>(s + o_pfe) : nullptr;
  h->cap_stream = nullptr;
  int rc = t3_prepare_kernel(t3_megakernel<__nv_bfloat16>);
  if (!rc) rc = t3_prepare_kernel(t3_megakernel<float>);
  if (rc) { cudaFree(h->slab); delete h; return CBX_ERR; }
  *out = h;
  return CBX_OK;
}

extern "C" int cbx_t3_destroy(cbx_t3* h) {
  if (!h) return CBX_OK;
  for (auto& g : h->pf_graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  cudaFree(h->slab);
  delete h;
  return CBX_OK;
}

static int t3_launch(cbx_t3* h, T3Params& p, cudaStream_t st) {
  CBX_CUDA(cudaMemsetAsync(p.ll, 0, LL_WORDS * 8, st));     // flags restart at 1 in every launch
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
  CBX_ON_DEVICE(h->cfg.device);
  T3Params p = h->base;
  h->B = B; h->host_pos = P; h->host_gen = 0;
  // reset per-utterance state: cache length 0, generated set = {BOS 6561} (t3.py:298)
  CBX_CUDA(cudaMemsetAsync(p.st, 0, sizeof(T3State), st));
  CBX_CUDA(cudaMemsetAsync(p.seen, 0, VOCAB_PAD, st));
  CBX_CUDA(cudaMemsetAsync(p.seen + 6561, 1, 1, st));
  p.mode = 0; p.B = B; p.embeds = embeds; p.P = P;
  if (h->al_layer >= 0) {
    h->al_rows0 = P - (h->al_i + h->al_n);        // query rows of the prompt behind the text: BOS (+ the doubled BOS)
    CBX_CHECK(h->al_rows0 >= 1 && h->al_rows0 <= 2, "cbx_t3_prefill: alignment analyzer: %d prompt rows after the text slice [%d, %d) "
              "(expected 1 or 2: P = %d)", h->al_rows0, h->al_i, h->al_i + h->al_n, P);
    static const AlState al_init = {0, 0, 0, 0, 0, -INFINITY, -INFINITY, -INFINITY, 0.f, 0.f, 0.f, 0.f, 0, 0, 0, 0};
    CBX_CUDA(cudaMemcpyAsync(h->al_state, &al_init, sizeof(AlState), cudaMemcpyHostToDevice, st));
  }
  static const bool force_mega = [] { const char* v = getenv("CBX_T3_PREFILL"); return v && v[0] == 'm'; }();
  if (h->prefill_ws && !force_mega) {
    // the prompt through each weight matrix once (tcgen05 GEMMs); leaves the slab, the position counter and the
    // pending logits as the megakernel's own prefill mode would
    cbx::T3PrefillArgs a{};
    a.embeds = embeds; a.B = B; a.P = P; a.n_layers = h->cfg.n_layers; a.max_pos = h->cfg.max_pos; a.kv_fp32 = h->cfg.kv_fp32;
    a.wplain = h->w.wplain; a.ln1 = p.ln1; a.ln2 = p.ln2; a.final_norm = p.final_norm;
    a.rope_cos = p.rope_cos; a.rope_sin = p.rope_sin; a.kcache = p.kcache; a.vcache = p.vcache;
    a.logits = p.logits; a.state = reinterpret_cast<int*>(p.st); a.workspace = h->prefill_ws;
    a.al_layer = h->al_layer; a.al_i = h->al_i; a.al_n = h->al_n; a.al_rows0 = h->al_rows0; a.al_out = h->al_prefill; a.al_cap = AL_CAP;
    h->launches += 1 + h->cfg.n_layers * 8 + 2;
    static const bool use_graph = [] { const char* v = getenv("CBX_T3_PREFILL_GRAPH"); return !(v && v[0] == '0'); }();
    cbx_t3::PrefillGraph* g = nullptr;
    if (use_graph && h->al_layer < 0) {          // analyzer runs prefill eagerly (its extra kernel is not part of the cached graphs)
      for (auto& c : h->pf_graphs) if (c.B == B && c.P == P) { g = &c; break; }
      if (!g) {
        if (h->pf_graphs.size() >= 32) {            // bound the cache: drop the oldest entry
          if (h->pf_graphs.front().exec) cudaGraphExecDestroy(h->pf_graphs.front().exec);
          h->pf_graphs.erase(h->pf_graphs.begin());
        }
        h->pf_graphs.push_back({B, P, 0, nullptr});
        g = &h->pf_graphs.back();
      }
      g->seen += 1;
    }
    if (!g || g->seen < 2) return cbx::t3_prefill_gemm(a, st);          // eager
    if (!g->exec) {
      if (!h->cap_stream) CBX_CUDA(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
      cbx::T3PrefillArgs ag = a;
      ag.embeds = h->pf_embeds;
      cudaGraph_t graph = nullptr;
      CBX_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
      const int rc = cbx::t3_prefill_gemm(ag, h->cap_stream);
      const cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
      if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cbx_set_error("cbx_t3_prefill: graph capture failed (%s)", cudaGetErrorString(ce));
        return CBX_ERR;
      }
      const cudaError_t ie = cudaGraphInstantiate(&g->exec, graph, 0);
      cudaGraphDestroy(graph);
      CBX_CHECK(ie == cudaSuccess, "cbx_t3_prefill: cudaGraphInstantiate failed (%s)", cudaGetErrorString(ie));
    }
    CBX_CUDA(cudaMemcpyAsync(h->pf_embeds, embeds, (size_t)B * P * T3_H * sizeof(float), cudaMemcpyDeviceToDevice, st));
    CBX_CUDA(cudaGraphLaunch(g->exec, st));
    return CBX_OK;
  }
  CBX_CHECK(h->al_layer < 0, "cbx_t3_prefill: the alignment analyzer needs the GEMM prefill (CbxT3Weights.wplain)");
  return t3_launch(h, p, st);
}

extern "C" int cbx_t3_decode(cbx_t3* h, const CbxSampleParams* sp, int max_new_tokens, const int32_t* forced_tokens,
                             int32_t* tokens_out, int32_t* n_out, float* logits_out, void* stream) {
  CBX_CHECK(h && sp && tokens_out && n_out, "cbx_t3_decode: null argument");
  CBX_CHECK(max_new_tokens >= 1, "cbx_t3_decode: max_new_tokens must be >= 1");
  CBX_CHECK(sp->temperature > 0.f, "cbx_t3_decode: temperature must be > 0");
  CBX_CHECK(sp->repetition_penalty > 0.f, "cbx_t3_decode: repetition_penalty must be > 0");
  CBX_CHECK(h->host_pos > 0, "cbx_t3_decode: no prefill on this handle yet");
  CBX_CHECK(h->host_pos + max_new_tokens <= h->cfg.max_pos, "cbx_t3_decode: %d cached + %d new positions exceed the KV slab (max_pos %d)",
            h->host_pos, max_new_tokens, h->cfg.max_pos);
  CBX_CHECK(h->cfg.speech_pos_rows <= 0 || h->host_gen + max_new_tokens + 1 <= h->cfg.speech_pos_rows,
            "cbx_t3_decode: %d + %d tokens exceed the learned speech position table (%d rows)", h->host_gen, max_new_tokens,
            h->cfg.speech_pos_rows);
  h->host_pos += max_new_tokens; h->host_gen += max_new_tokens;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CBX_ON_DEVICE(h->cfg.device);
  T3Params p = h->base;
  p.mode = 1; p.B = h->B; p.sp = *sp; p.max_new = max_new_tokens; p.forced = forced_tokens;
  p.tokens_out = tokens_out; p.n_out = n_out; p.logits_out = logits_out;
  p.prof = h->prof; p.prof_it = h->prof_it;
  if (h->al_layer >= 0) {
    CBX_CHECK(h->host_gen == max_new_tokens, "cbx_t3_decode: with the alignment analyzer the decode launch must follow the prefill directly "
              "(the attention row of an earlier launch's last step is gone)");
    p.al_layer = h->al_layer; p.al_i = h->al_i; p.al_n = h->al_n; p.al_rows0 = h->al_rows0;
    p.al_prefill = h->al_prefill; p.al_inject = h->al_inject; p.al_rows_out = h->al_rows_out; p.al_out = h->al_out;
    p.al_state = h->al_state;
  }
  return t3_launch(h, p, st);
}

extern "C" int cbx_t3_set_alignment(cbx_t3* h, int layer, int text_begin, int n_text, const float* inject_rows, float* rows_out,
                                    int32_t* decisions_out) {
  CBX_CHECK(h, "cbx_t3_set_alignment: null handle");
  if (layer < 0) { h->al_layer = -1; return CBX_OK; }
  CBX_CHECK(layer < h->cfg.n_layers, "cbx_t3_set_alignment: layer %d >= n_layers %d", layer, h->cfg.n_layers);
  CBX_CHECK(!h->cfg.kv_fp32, "cbx_t3_set_alignment: the analyzer reads the tile attention path (bf16 KV slab only)");
  CBX_CHECK(h->cfg.max_pos <= MAXS * ATT_TP, "cbx_t3_set_alignment: max_pos %d exceeds the tile attention path (%d positions)",
            h->cfg.max_pos, MAXS * ATT_TP);
  CBX_CHECK(h->prefill_ws, "cbx_t3_set_alignment: needs the GEMM prefill (CbxT3Weights.wplain)");
  CBX_CHECK(n_text >= 6 && n_text <= AL_CAP && text_begin >= 0, "cbx_t3_set_alignment: text slice [%d, %d + %d) unsupported (6..%d columns)",
            text_begin, text_begin, n_text, AL_CAP);
  h->al_layer = layer; h->al_i = text_begin; h->al_n = n_text;
  h->al_inject = inject_rows; h->al_rows_out = rows_out; h->al_out = decisions_out;
  return CBX_OK;
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
