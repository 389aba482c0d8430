// tcgen05 flash attention for the S3Gen estimator's BasicTransformerBlock (reference: matcha/transformer.py:196-204 ->
// diffusers Attention / AttnProcessor2_0: 8 heads x 64, scale 1/8, bidirectional, key-padding mask only
// (decoder.py:25-33, utils/mask.py:167-193)).
//
// One CTA owns 128 query rows of one (batch row, head).  Keys are taken in blocks of 128:
//   warp 8            the issuer: one thread issues every TMA load (Q tile once; K blocks [128 keys][64 d] into 4 slots,
//                     V blocks into 3 slots, 128-byte swizzle, straight out of the q|k|v buffer the GEMM wrote) and every
//                     tcgen05.mma, in the order their preconditions become true:
//                       S_j = Q K_j^T    M128 N128 K16 x 4, both operands K-major -> TMEM S buffer j % 3, issued three
//                                        blocks ahead, as soon as the buffer's previous reader is done
//                       O~_j = P_j V_j   M128 N64 K16 x 8; A = P from shared memory, B = the V block as it lies in
//                                        memory ([keys][d], d contiguous) = an MN-major operand: no transpose -> TMEM O[g]
//   warps 0-3 / 4-7   two softmax groups taking alternate key blocks (g = block & 1), thread <-> query row (TMEM lane).
//                     Default (ATC_ONEPASS): ONE pass over the S tile — the row's 128 scores sit in registers between the
//                     maximum and the exponentials, bf16 P is written back over the S columns in TMEM (A operand of a TS-mode
//                     tcgen05.mma), the group's running output accumulates in TMEM across its key blocks and is rescaled
//                     there only when the row's reference maximum has to be raised (see the loop).  -DATC_ONEPASS=0 / 1
//                     keeps the first form for this kernel: two passes (row max; exp2 / row sum / P), online rescaling
//                     in registers with O~ read back from TMEM one block late.
// The two groups' (m, l, o) states are merged through shared memory at the end (as the mma.sync kernel did).
// fp32 scores, fp32 online softmax, bf16 P (as before).  Replaces the mma.sync kernel of csrc/s3_attn.cu (kept: CBX_ATTN_TC=0).
// (A variant in which thread 0 of each softmax group issued its own loads and MMAs — no issuer warp, 255 registers — was
// slower: the issue sequence sits on the group's critical path.  33.6 vs 32.6 ms per solve.)
#include "s3_gemm.cuh"
#include <cstdio>
#include <cstdlib>

namespace cbx {
namespace {

#ifndef ATC_P_TMEM
#define ATC_P_TMEM 1          // P travels to the second product through TMEM (in place over the S tile); 0: through shared memory
#endif
#ifndef ATC_ONEPASS
#define ATC_ONEPASS 3         // bit (NG - 1): that instantiation reads each S tile from TMEM ONCE (see the softmax loop below)
#endif
constexpr int TQ = 128, TK = 128, TD = 64;
constexpr int TILE_B = TQ * TD * 2;                    // 16 KB: [128][64] bf16
constexpr int P_B = TQ * TK * 2;                       // 32 KB: two k-blocks of [128][64]
// NG = softmax groups per CTA.  2 (one utterance: 112 CTAs, one per SM): the groups take alternate key blocks, 3 S tiles
// in TMEM, 4 K / 3 V slots, 512 TMEM columns.  1 (padded batches: thousands of CTAs): one group walks all key blocks,
// 1 S tile, 2 K / 2 V slots, 256 TMEM columns, 80 KB of shared memory — TWO CTAs per SM overlap each other's latencies
// instead of two groups inside one CTA, and there is nothing to merge.
template <int NG> struct AtcCfg {
  static constexpr int NSK = NG == 2 ? 4 : 2, NST = NG == 2 ? 3 : 2, NSB = NG == 2 ? 3 : 1;
  static constexpr int OFF_Q = 0, OFF_K = TILE_B, OFF_V = OFF_K + NSK * TILE_B, OFF_P = OFF_V + NST * TILE_B;
  static constexpr int OFF_ML = OFF_P + (NG == 2 ? 2 * P_B : 0), OFF_BAR = OFF_ML + (NG == 2 ? TQ * 2 * 4 : 0);
  static constexpr int THREADS = 32 * (4 * NG + 1);      // softmax warps + the issuer warp
  static constexpr int SMEM = OFF_BAR + 512 + 1024;
  static constexpr int TMEM_COLS = NG == 2 ? 512 : 256;
};
static_assert(TQ * TD * 4 <= P_B, "group-1 output state fits its P buffer");
static_assert(ATC_P_TMEM || true, "");

struct AttnTcParams {
  CUtensorMap tm;            // q|k|v bf16 [B][L][1536], box {64, 128, 1}
  __nv_bfloat16* out;        // [B][L][512]
  const int* lens;
  int L;
  int nx, n_items;           // persistent two-tile kernel: query-tile pairs per (batch row, head); work items = B x 8 x nx
  unsigned long long* dbg;   // optional clock64 stamps of CTA (0, 0, 0), 32 slots per launch (tools/attn_stamps.py, attn_batch_stamps.py)
};
// stamps are SM cycle counts (clock64): every stamp of a launch comes from CTA (0, 0, 0), i.e. one SM — finer than %globaltimer
__device__ __forceinline__ unsigned long long atc_gtimer() { return (unsigned long long)clock64(); }
#define ATC_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) p.dbg[k] = atc_gtimer(); } while (0)

__device__ __forceinline__ float ex2f(float x) {
#ifdef ATC_ABL_NOEX2
  return x * 0.001f + 1.0f;
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
  unsigned spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();
  }
}
template <int ID> __device__ __forceinline__ void group_bar() { asm volatile("bar.sync %0, 128;" ::"n"(ID) : "memory"); }

// TMEM -> registers without the wait: the caller overlaps the load with arithmetic on the previous chunk
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM: this warp's 32 lanes x 16 consecutive 32-bit columns
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]: the A operand (128 lanes x 8 columns of packed bf16 pairs per K = 16 step) comes from TMEM
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int NG>
__global__ void __launch_bounds__(AtcCfg<NG>::THREADS, NG == 1 ? 2 : 1) s3_attn_tc_kernel(const __grid_constant__ AttnTcParams p) {
  using C = AtcCfg<NG>;
  constexpr int NSK = C::NSK, NST = C::NST, NSB = C::NSB, OFF_Q = C::OFF_Q, OFF_K = C::OFF_K, OFF_V = C::OFF_V, OFF_P = C::OFF_P,
                OFF_ML = C::OFF_ML, OFF_BAR = C::OFF_BAR;
  static_assert(NG == 2 || ATC_P_TMEM, "the one-group form keeps P in TMEM (no shared-memory P buffers)");
  constexpr bool ONEPASS = ((ATC_ONEPASS >> (NG - 1)) & 1) != 0;
  static_assert(!ONEPASS || ATC_P_TMEM, "the one-pass softmax keeps P in TMEM");
  // one group, one pass: P gets its OWN 64 TMEM columns (S [0, 128) | P [128, 192) | O~ [192, 256)), so the next block's
  // Q K^T may overwrite the S tile as soon as the four softmax warps hold the row in registers (s_free) — it runs under
  // the exponentials of this block instead of after this block's P V product
  constexpr bool SPLIT_P = ONEPASS && NG == 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem + OFF_Q;
  uint8_t* sK = smem + OFF_K;
  uint8_t* sV = smem + OFF_V;
  uint8_t* sP = smem + OFF_P;
  float* sML = reinterpret_cast<float*>(smem + OFF_ML);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_q = bars;
  uint64_t* k_full = bars + 1;           // [NSK] K block landed
  uint64_t* v_full = k_full + NSK;       // [NST] V block landed
  uint64_t* v_free = v_full + NST;       // [NST] its P V product has completed
  uint64_t* s_full = v_free + NST;       // [NSB] S tile of block j in TMEM buffer j % NSB
  uint64_t* p_full = s_full + NSB;       // [2]   P of the group's block written, S buffer read (4 arrivals: one per warp)
  uint64_t* o_full = p_full + 2;         // [2]   the group's P V product has completed
  uint64_t* s_free = o_full + 2;         // [1]   SPLIT_P: the S tile has been read into registers (4 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * TQ, head = blockIdx.y, b = blockIdx.z;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    ATC_STAMP(0);
    mbar_init(bar_q, 1);
    for (int i = 0; i < NSK; ++i) mbar_init(&k_full[i], 1);
    for (int i = 0; i < NST; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_free[i], 1); }
    for (int i = 0; i < NSB; ++i) mbar_init(&s_full[i], 1);
    for (int i = 0; i < NG; ++i) { mbar_init(&p_full[i], 4); mbar_init(&o_full[i], 1); }
    mbar_init(s_free, 4);
    mbar_fence_init();
    tma_prefetch_desc(&p.tm);
  }
  if (warp == 0) tmem_alloc(tmem_slot, C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_o = tmem_base + NSB * TK + (SPLIT_P ? TK / 2 : 0);      // O~[NG] behind the S buffers (and the P tile)
  const int Lk = p.lens ? min(max(p.lens[b], 1), p.L) : p.L;
  const int nkb = (Lk + TK - 1) / TK;

  if (warp == 4 * NG) {
    // ---- the issuer: every TMA load and every tcgen05.mma of the CTA, in the order their preconditions become true ----
    if (elect_one()) {     // elect.sync, not lane == 0: the compiler then passes descriptors in uniform registers (tc.cuh)
      const uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
      const uint32_t idesc_o = umma_idesc_bf16(TQ, TD) | (1u << 16);        // B (= V) is MN-major
      auto load_k = [&](int j) {
        const int s = j % NSK;
        mbar_arrive_expect_tx(&k_full[s], TILE_B);
        tma_load_3d(sK + s * TILE_B, &p.tm, &k_full[s], 512 + head * TD, j * TK, b);
      };
      auto load_v = [&](int j) {
        const int s = j % NST;
        mbar_arrive_expect_tx(&v_full[s], TILE_B);
        tma_load_3d(sV + s * TILE_B, &p.tm, &v_full[s], 1024 + head * TD, j * TK, b);
      };
      const uint64_t dq = umma_desc_k128(smem_u32(sQ));
      auto issue_s = [&](int j) {            // S_j = Q K_j^T into S buffer j % NSB (the caller knows the buffer is free)
        const int s = j % NSK, sb = j % NSB;
        mbar_wait_sleep(&k_full[s], (j / NSK) & 1);
        tc_fence_after();
        const uint64_t db = umma_desc_k128(smem_u32(sK + s * TILE_B));
#pragma unroll
        for (int k = 0; k < TD / 16; ++k) umma_bf16(tmem_base + sb * TK, dq + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_s, k != 0);
        umma_commit(&s_full[sb]);
      };
      pdl_wait();                                  // q|k|v come from the GEMM just before
      ATC_STAMP(1);
      mbar_arrive_expect_tx(bar_q, TILE_B);
      tma_load_3d(sQ, &p.tm, bar_q, head * TD, q0, b);
      for (int j = 0; j < NSK && j < nkb; ++j) { load_k(j); if (j < NST) load_v(j); }
      mbar_wait_sleep(bar_q, 0);
      for (int j = 0; j < NSB && j < nkb; ++j) issue_s(j);
      if constexpr (SPLIT_P) {
        for (int j = 0; j < nkb; ++j) {
          const int s = j % NST;
          mbar_wait_sleep(s_free, j & 1);            // S_j is in registers: the S tile and K slot j % NSK are free
          if (j + NSK < nkb) load_k(j + NSK);
          if (j + 1 < nkb) issue_s(j + 1);           // runs while the softmax warps exponentiate block j
          mbar_wait_sleep(&p_full[0], j & 1);
          if (j >= 1 && j - 1 + NST < nkb) {
            mbar_wait_sleep(&v_free[(j - 1) % NST], ((j - 1) / NST) & 1);
            load_v(j - 1 + NST);
          }
          mbar_wait_sleep(&v_full[s], (j / NST) & 1);
          tc_fence_after();
          const uint32_t va = smem_u32(sV + s * TILE_B);
#pragma unroll
          for (int k = 0; k < TK / 16; ++k)
            umma_bf16_ts(tmem_o, tmem_base + TK + k * 8, umma_desc_k128(va + k * 2048), idesc_o, (k != 0) || (j > 0));
          umma_commit(&o_full[0]);
          umma_commit(&v_free[s]);
        }
      } else
      for (int j = 0; j < nkb; ++j) {
        const int g = j % NG, it = j / NG, s = j % NST;
        // P_j is in shared memory, and its group is done with S buffer j % NSB (so S_j has long completed: K slot free)
        mbar_wait_sleep(&p_full[g], it & 1);
        if (j + NSK < nkb) load_k(j + NSK);
        if (j >= 1 && j - 1 + NST < nkb) {       // V slot of block j - 1: its product was issued one iteration ago
          mbar_wait_sleep(&v_free[(j - 1) % NST], ((j - 1) / NST) & 1);
          load_v(j - 1 + NST);
        }
        mbar_wait_sleep(&v_full[s], (j / NST) & 1);
        tc_fence_after();
        [[maybe_unused]] const uint32_t pa = smem_u32(sP + g * P_B);
        const uint32_t va = smem_u32(sV + s * TILE_B);
#pragma unroll
        for (int k = 0; k < TK / 16; ++k) {
          // A: k-block k / 4 of P, 32 bytes per K = 16 step inside the swizzle row; B: 16 key rows = 2048 bytes per step
          const uint64_t db = umma_desc_k128(va + k * 2048);
#if ATC_P_TMEM
          umma_bf16_ts(tmem_o + g * TD, tmem_base + (j % NSB) * TK + k * 8, db, idesc_o, (k != 0) || (ONEPASS && it > 0));
#else
          const uint64_t da = umma_desc_k128(pa + (k >> 2) * TILE_B) + (uint64_t)((k & 3) * 2);
          umma_bf16(tmem_o + g * TD, da, db, idesc_o, k != 0);
#endif
        }
        umma_commit(&o_full[g]);
        umma_commit(&v_free[s]);
        if (j + NSB < nkb) issue_s(j + NSB);       // into the S buffer block j's group just finished with
      }
    }
    __syncwarp();
  } else {
  const int g = warp >> 2, q = warp & 3, row = q * 32 + lane;
  const uint32_t lane_off = (uint32_t)(q * 32) << 16;
  const uint32_t tO = tmem_o + lane_off + g * TD;
  const float sc = 0.125f * 1.4426950408889634f;              // 1/sqrt(64) * log2(e): scores in the log2 domain
  float m = -INFINITY, l = 0.f;
  [[maybe_unused]] float corr_prev = 1.f;
  float o[TD];
#pragma unroll
  for (int d = 0; d < TD; ++d) o[d] = 0.f;
  [[maybe_unused]] uint8_t* prow = sP + g * P_B + row * 128;
  int i = 0;
  if constexpr (ONEPASS) {
    // One read of the S tile: the whole row (128 scores) sits in registers between the maximum and the exponentials, and
    // the running output stays in TMEM — the P V products of consecutive blocks ACCUMULATE there, nothing is folded in
    // registers.  The reference maximum m of a row is only raised when a block's maximum exceeds it by more than 8 (log2
    // domain: p <= 256, far inside bf16 / fp32 range); then, and only then, the row's O~ and l are rescaled by
    // exp2(m_old - m_new).  O / l does not depend on the reference, so the result is the same softmax.  Per key block
    // the TMEM traffic of a row drops from 2 x 128 (S twice) + 64 (O~) reads to 128.
    for (int j = g; j < nkb; j += NG, ++i) {
      const int sb = j % NSB;
      const uint32_t tS = tmem_base + lane_off + sb * TK;
      mbar_wait_sleep(&s_full[sb], (j / NSB) & 1);
      tc_fence_after();
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(2);
      if (threadIdx.x == 0 && i == 3) ATC_STAMP(10);
      const int kbase = j * TK;
      uint32_t s[4][32];
#ifdef ATC_ABL_LD1            // timing-only ablation: a quarter of the TMEM reads
      tmem_ld32_issue(tS, s[0]);
      tmem_ld_wait();
#pragma unroll
      for (int c = 1; c < 4; ++c)
#pragma unroll
        for (int k = 0; k < 32; ++k) s[c][k] = s[0][k] ^ (uint32_t)(c * 4096);
#else
#pragma unroll
      for (int c = 0; c < 4; ++c) tmem_ld32_issue(tS + c * 32, s[c]);
      tmem_ld_wait();
#endif
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(4);
      if constexpr (SPLIT_P) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);
      }
      const uint32_t tP = SPLIT_P ? tmem_base + lane_off + TK : tS;
      if (kbase + TK > Lk) {       // keys past the row's length: -inf (exp2 -> 0)
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int k = 0; k < 32; ++k)
            if (kbase + c * 32 + k >= Lk) s[c][k] = 0xff800000u;
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int k = 0; k < 32; ++k) mx4[k & 3] = fmaxf(mx4[k & 3], __uint_as_float(s[c][k]));
      const float mb = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * sc;
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(3);
      // Every phase of o_full is observed, in order (a parity wait cannot tell phases two apart): the group's previous
      // product — the last writer of O~, and with SPLIT_P the last reader of the P tile — has completed.  Needed before a
      // rescale, else before the first P store; it was issued a whole softmax ago, so this rarely spins.
      bool waited = i == 0;
      auto wait_prev = [&]() {
        if (!waited) { mbar_wait_sleep(&o_full[g], (i - 1) & 1); tc_fence_after(); waited = true; }
      };
      const bool need = mb > m + 8.0f;              // first block: m = -inf
      if (__any_sync(0xffffffffu, need)) {          // warp-uniform: tcgen05.ld / .st are warp-collective
        const float m_new = need ? mb : m;
        if (i > 0) {
          const float f = ex2f(m - m_new);          // 1 for the rows of the warp that keep their reference
          wait_prev();
#pragma unroll
          for (int c = 0; c < 4; ++c) {             // 16 columns at a time: the score row stays in registers beside it
            uint32_t r[16];
            tmem_ld16_issue(tO + c * 16, r);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 16; ++k) r[k] = __float_as_uint(__uint_as_float(r[k]) * f);
            tmem_st16(tO + c * 16, r);
          }
          l *= f;
        }
        m = m_new;
      }
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(5);
      float ls4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pw[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float p0 = ex2f(fmaf(__uint_as_float(s[c][2 * e]), sc, -m)), p1 = ex2f(fmaf(__uint_as_float(s[c][2 * e + 1]), sc, -m));
#ifdef ATC_ABL_TRUNC          // timing-only ablation: one byte-permute instead of the rounding conversion
          pw[e] = __byte_perm(__float_as_uint(p0), __float_as_uint(p1), 0x7632);
#else
          pw[e] = pack_bf16x2(p0, p1);
#endif
          // the row sum is taken over the ROUNDED probabilities, the values the tensor core multiplies: with a lazily
          // raised reference the dominant key's p is no longer exactly 1, and its bf16 rounding would otherwise appear
          // in the output unnormalised (numerator rounded, denominator not)
          ls4[e & 3] += __uint_as_float(pw[e] << 16) + __uint_as_float(pw[e] & 0xffff0000u);
        }
        if (c == 0) { if (threadIdx.x == 0 && i == 2) ATC_STAMP(26); wait_prev(); if (threadIdx.x == 0 && i == 2) ATC_STAMP(27); }
        tmem_st16(tP + c * 16, pw);          // P chunk c (over S columns [16 c, 16 c + 16) unless SPLIT_P: the row is in registers)
      }
      l += (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(6);
      tmem_st_wait();
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(7);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[g]);
      if (threadIdx.x == 0 && i == 2) ATC_STAMP(8);
    }
    if (i > 0) {
      mbar_wait_sleep(&o_full[g], (i - 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld32(tO + c * 32, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) o[c * 32 + k] = v[k];
      }
    }
  } else {
  for (int j = g; j < nkb; j += NG, ++i) {
    const int sb = j % NSB;
    const uint32_t tS = tmem_base + lane_off + sb * TK;
    mbar_wait_sleep(&s_full[sb], (j / NSB) & 1);
    tc_fence_after();
    if (threadIdx.x == 0 && i == 1) ATC_STAMP(2);
    if (threadIdx.x == 0 && i == 2) ATC_STAMP(10);
    if (threadIdx.x == 0 && i == 3) ATC_STAMP(11);
    const int kbase = j * TK;
    const bool tail = kbase + TK > Lk;
    // pass 1: row maximum, 32-column chunks (the load of chunk c + 1 is in flight while chunk c is reduced)
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // four independent chains (2 warps per scheduler:
    {                                                                  // a 128-long dependent chain would be exposed)
      uint32_t ra[32], rb[32];
      auto red = [&](const uint32_t (&r)[32], int c) {
        if (tail) {
#pragma unroll
          for (int k = 0; k < 32; ++k) mx4[k & 3] = fmaxf(mx4[k & 3], (kbase + c * 32 + k >= Lk) ? -INFINITY : __uint_as_float(r[k]));
        } else {
#pragma unroll
          for (int k = 0; k < 32; ++k) mx4[k & 3] = fmaxf(mx4[k & 3], __uint_as_float(r[k]));
        }
      };
#ifndef ATC_ABL_NOPASS1      // timing-only ablation builds (results are garbage): -DATC_ABL_NOPASS1 / -DATC_ABL_NOEX2 / -DATC_ABL_NOFOLD
      tmem_ld32_issue(tS, ra);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 32, rb);
      red(ra, 0);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 64, ra);
      red(rb, 1);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 96, rb);
      red(ra, 2);
      tmem_ld_wait();
      red(rb, 3);
#else
      mx4[0] = 8.f;
#endif
    }
    const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
    if (threadIdx.x == 0 && i == 1) ATC_STAMP(3);
    const float m_new = fmaxf(m, mx * sc);
    const float corr = ex2f(m - m_new);           // first block: exp2(-inf) = 0, and o = l = 0
    // fold the previous block's P V product into the running output (it has been computing meanwhile); its completion
    // also frees the group's P buffer
    if (i > 0) {
      mbar_wait_sleep(&o_full[g], (i - 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 0 && i == 1) ATC_STAMP(4);
#ifndef ATC_ABL_NOFOLD
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld32_issue(tO + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 32; ++k) o[c * 32 + k] = fmaf(o[c * 32 + k], corr_prev, __uint_as_float(r[k]));
      }
#endif
    }
    corr_prev = corr;
    if (threadIdx.x == 0 && i == 1) ATC_STAMP(5);
    // pass 2: p = exp2(s - m), row sum, bf16 P in the swizzled K-major layout
    float ls4[4] = {0.f, 0.f, 0.f, 0.f};
    {
      uint32_t ra[32], rb[32];
      auto emit = [&](uint32_t (&r)[32], int c) {
        [[maybe_unused]] uint8_t* pk = prow + (c >> 1) * TILE_B;
        [[maybe_unused]] uint32_t pw[16];
        if (tail) {          // keys past the row's length: exp2(-inf) = 0
#pragma unroll
          for (int k = 0; k < 32; ++k) if (kbase + c * 32 + k >= Lk) r[k] = 0xff800000u;
        }
#pragma unroll
        for (int jc = 0; jc < 4; ++jc) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int k = jc * 8 + e * 2;
            const float p0 = ex2f(fmaf(__uint_as_float(r[k]), sc, -m_new)), p1 = ex2f(fmaf(__uint_as_float(r[k + 1]), sc, -m_new));
            ls4[e] += p0 + p1;
            w[e] = pack_bf16x2(p0, p1);
          }
#if ATC_P_TMEM
#pragma unroll
          for (int e = 0; e < 4; ++e) pw[jc * 4 + e] = w[e];
#else
          *reinterpret_cast<uint4*>(pk + ((((c & 1) * 4 + jc) ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
#endif
        }
#if ATC_P_TMEM
        tmem_st16(tS + c * 16, pw);          // P chunk c over S columns [16 c, 16 c + 16): all of them already read
#endif
      };
      tmem_ld32_issue(tS, ra);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 32, rb);
      emit(ra, 0);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 64, ra);
      emit(rb, 1);
      tmem_ld_wait();
      tmem_ld32_issue(tS + 96, rb);
      emit(ra, 2);
      tmem_ld_wait();
      emit(rb, 3);
    }
    if (threadIdx.x == 0 && i == 1) { if (ls4[0] == 123.456f) ATC_STAMP(12); ATC_STAMP(6); }
    l = l * corr + ((ls4[0] + ls4[1]) + (ls4[2] + ls4[3]));
    m = m_new;
#if ATC_P_TMEM
    tmem_st_wait();
#else
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // P visible to the tensor core
#endif
    if (threadIdx.x == 0 && i == 1) ATC_STAMP(7);
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&p_full[g]);    // one arrive per warp: 32 P rows written, and the warp is done with the S buffer
    if (threadIdx.x == 0 && i == 1) ATC_STAMP(8);
  }
  if (i > 0) {
    mbar_wait_sleep(&o_full[g], (i - 1) & 1);
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float v[32];
      tmem_ld32(tO + c * 32, v);
#pragma unroll
      for (int k = 0; k < 32; ++k) o[c * 32 + k] = fmaf(o[c * 32 + k], corr_prev, v[k]);
    }
  }
  }
  tc_fence_before();
  if (threadIdx.x == 0) ATC_STAMP(14);
  // ---- merge group 1 into group 0 (group 1's P buffer is idle: its last product has completed); one group: nothing to merge ----
  float c0 = 1.f, c1 = 0.f, l_all = l;
  [[maybe_unused]] float* xo = reinterpret_cast<float*>(sP + P_B);           // [64 d][128 rows]
  if constexpr (NG == 2) {
    if (g == 1) {
#pragma unroll
      for (int d = 0; d < TD; ++d) xo[d * TQ + row] = o[d];
      sML[row * 2] = m; sML[row * 2 + 1] = l;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (g == 0) {
      const float m1 = sML[row * 2], l1 = sML[row * 2 + 1];
      const float mn = fmaxf(m, m1);
      c0 = (m == -INFINITY) ? 0.f : ex2f(m - mn); c1 = (m1 == -INFINITY) ? 0.f : ex2f(m1 - mn);
      l_all = l * c0 + l1 * c1;
    }
  }
  if (g == 0) {
    const float inv = 1.0f / l_all;
    const int t = q0 + row;
    if (t < p.L) {
      __nv_bfloat16* dst = p.out + ((size_t)b * p.L + t) * 512 + head * TD;
#pragma unroll
      for (int d8 = 0; d8 < TD / 8; ++d8) {
        float y[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float v = o[d8 * 8 + e] * c0;
          if constexpr (NG == 2) v += xo[(d8 * 8 + e) * TQ + row] * c1;
          y[e] = v * inv;
        }
        *reinterpret_cast<uint4*>(dst + d8 * 8) = make_uint4(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]),
                                                             pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
      }
    }
  }
  }
  __syncthreads();
  if (threadIdx.x == 0) ATC_STAMP(15);
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// Padded batches: a PERSISTENT kernel, TWO query tiles of one (batch row, head) per work item.
//   * warps 0-3 own query tile 2 x, warps 4-7 tile 2 x + 1: two independent one-pass softmax pipelines (the loop body is
//     the one above: row in registers, O~ accumulating in TMEM, lazily raised reference, P in its own TMEM columns), two
//     warps per SM sub-partition.  They share every K and V block: one TMA load feeds both tiles' products.
//   * warp 8 issues the score products and owns the Q buffers (two: the next item's Q lands during this item) and the K
//     ring; warp 9 issues the P V products and owns the V ring.  Neither ever waits for something the other side needs.
//   * one CTA per SM walks work items blockIdx.x, + gridDim.x, ...: the K / V rings, the barriers' phases and TMEM run on
//     ACROSS items, so an item's first blocks are in shared memory — and its first score product issued — before the
//     previous item's last block is finished.  The one-item-per-CTA form paid ~2.5 us of load latency plus prologue and
//     epilogue per 16 us item with nothing to overlap it (512 TMEM columns: one CTA per SM).
//   TMEM, per tile g: S [256 g, +128) | P [256 g + 128, +64) | O~ [256 g + 192, +64)
struct QtCfg {
  static constexpr int NSK = 4, NST = 4;
  static constexpr int OFF_Q = 0, OFF_K = 4 * TILE_B, OFF_V = OFF_K + NSK * TILE_B, OFF_BAR = OFF_V + NST * TILE_B;
  static constexpr int THREADS = 32 * 10, SMEM = OFF_BAR + 512 + 1024;
};

struct QtItem {                 // one (batch row, head, query-tile pair)
  int w, b, head, q0, ntile, Lk, nkb;
  bool valid;
};
__device__ __forceinline__ QtItem qt_item(const AttnTcParams& p, int w) {
  QtItem it;
  it.w = w; it.valid = w < p.n_items;
  const int x = w % p.nx, t = w / p.nx;
  it.head = t & 7; it.b = t >> 3;
  it.q0 = x * 2 * TQ;
  it.ntile = it.q0 + TQ < p.L ? 2 : 1;              // an odd tile count leaves the last pair with one tile
  it.Lk = (it.valid && p.lens) ? min(max(p.lens[it.b], 1), p.L) : p.L;
  it.nkb = (it.Lk + TK - 1) / TK;
  return it;
}

__global__ void __launch_bounds__(QtCfg::THREADS, 1) s3_attn_qt_kernel(const __grid_constant__ AttnTcParams p) {
  constexpr int NSK = QtCfg::NSK, NST = QtCfg::NST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem + QtCfg::OFF_Q;     // [2 buffers][2 tiles]
  uint8_t* sK = smem + QtCfg::OFF_K;
  uint8_t* sV = smem + QtCfg::OFF_V;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + QtCfg::OFF_BAR);
  uint64_t* bar_q = bars;                // [2]   Q of item i in buffer i & 1
  uint64_t* k_full = bar_q + 2;          // [NSK] running K block n in slot n % NSK
  uint64_t* v_full = k_full + NSK;       // [NST]
  uint64_t* v_free = v_full + NST;       // [NST] the P V products of the block (both tiles) have completed
  uint64_t* s_full = v_free + NST;       // [2]   per tile: S of its next block is in TMEM
  uint64_t* s_free = s_full + 2;         // [2]   per tile: the S row is in registers (4 arrivals)
  uint64_t* p_full = s_free + 2;         // [2]   per tile: P written (4 arrivals)
  uint64_t* o_full = p_full + 2;         // [2]   per tile: the P V product has completed
  uint64_t* o_free = o_full + 2;         // [2]   per tile: the item's O~ has been read out (4 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stride = gridDim.x;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) mbar_init(&bar_q[i], 1);
    for (int i = 0; i < NSK; ++i) mbar_init(&k_full[i], 1);
    for (int i = 0; i < NST; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_free[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 4); mbar_init(&p_full[i], 4); mbar_init(&o_full[i], 1); mbar_init(&o_free[i], 4);
    }
    mbar_fence_init();
    tma_prefetch_desc(&p.tm);
  }
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ---- issuer 1: Q buffers, the K ring, S = Q K^T.  Every barrier phase is a running count across items. ----
    if (elect_one()) {
      const uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
      QtItem U = qt_item(p, blockIdx.x), LK = U;      // U: the item whose blocks are being consumed; LK: cursor of the K loads
      int lk_j = 0;
      uint32_t kload = 0, kuse = 0, item = 0, sfree_n[2] = {0, 0};
      auto k_prefetch = [&]() {                   // keep the ring full: running block kload goes to slot kload % NSK
        while (kload < kuse + NSK && LK.valid) {
          const int s = kload % NSK;
          mbar_arrive_expect_tx(&k_full[s], TILE_B);
          tma_load_3d(sK + s * TILE_B, &p.tm, &k_full[s], 512 + LK.head * TD, lk_j * TK, LK.b);
          ++kload;
          if (++lk_j == LK.nkb) { LK = qt_item(p, LK.w + stride); lk_j = 0; }
        }
      };
      auto load_q = [&](const QtItem& it, uint32_t buf) {
        mbar_arrive_expect_tx(&bar_q[buf], it.ntile * TILE_B);
        for (int g = 0; g < it.ntile; ++g) tma_load_3d(sQ + (buf * 2 + g) * TILE_B, &p.tm, &bar_q[buf], it.head * TD, it.q0 + g * TQ, it.b);
      };
      auto issue_s = [&](int g, uint32_t buf, uint32_t n) {       // S of tile g <- Q_g (buffer buf) x running K block n
        const int s = n % NSK;
        mbar_wait_sleep(&k_full[s], (n / NSK) & 1);
        tc_fence_after();
        const uint64_t dq = umma_desc_k128(smem_u32(sQ + (buf * 2 + g) * TILE_B)), db = umma_desc_k128(smem_u32(sK + s * TILE_B));
#pragma unroll
        for (int k = 0; k < TD / 16; ++k) umma_bf16(tmem_base + g * 256, dq + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_s, k != 0);
        umma_commit(&s_full[g]);
      };
      pdl_wait();                                  // q|k|v come from the GEMM just before
      if (U.valid) {
        load_q(U, 0);
        k_prefetch();
        mbar_wait_sleep(&bar_q[0], 0);
        for (int g = 0; g < U.ntile; ++g) issue_s(g, 0, 0);
      }
      while (U.valid) {
        const QtItem N = qt_item(p, U.w + stride);
        const uint32_t buf = item & 1, nbuf = buf ^ 1;
        if (N.valid) load_q(N, nbuf);              // that buffer's last reader (item - 1) has completed all its score products
        for (int j = 0; j < U.nkb; ++j) {
          // a tile's S row is in registers: its S columns are free — its next score product runs under this block's exponentials
          for (int g = 0; g < U.ntile; ++g) {
            mbar_wait_sleep(&s_free[g], sfree_n[g] & 1);
            ++sfree_n[g];
            if (j + 1 < U.nkb) issue_s(g, buf, kuse + 1);
          }
          if (j + 1 == U.nkb && N.valid) {         // last block: the next item's first score products, for the tiles it has
            mbar_wait_sleep(&bar_q[nbuf], ((item + 1) >> 1) & 1);
            for (int g = 0; g < N.ntile; ++g) issue_s(g, nbuf, kuse + 1);
          }
          ++kuse;                                  // every score product of the block has completed: its K slot is free
          k_prefetch();
        }
        U = N;
        ++item;
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ---- issuer 2: the V ring, O~ (+)= P V ----
    if (elect_one()) {
      const uint32_t idesc_o = umma_idesc_bf16(TQ, TD) | (1u << 16);        // B (= V) is MN-major
      QtItem U = qt_item(p, blockIdx.x), LV = U;
      int lv_j = 0;
      uint32_t vload = 0, n = 0, pfull_n[2] = {0, 0}, items_g[2] = {0, 0};
      auto load_v = [&]() {                        // running block vload -> slot vload % NST
        const int s = vload % NST;
        mbar_arrive_expect_tx(&v_full[s], TILE_B);
        tma_load_3d(sV + s * TILE_B, &p.tm, &v_full[s], 1024 + LV.head * TD, lv_j * TK, LV.b);
        ++vload;
        if (++lv_j == LV.nkb) { LV = qt_item(p, LV.w + stride); lv_j = 0; }
      };
      pdl_wait();
      while (vload < NST && LV.valid) load_v();
      while (U.valid) {
        for (int j = 0; j < U.nkb; ++j, ++n) {
          const int s = n % NST;
          if (n >= 1) {      // the block before: its products were issued one iteration ago; every v_free phase is observed, in order
            mbar_wait_sleep(&v_free[(n - 1) % NST], ((n - 1) / NST) & 1);
            if (LV.valid) load_v();                // vload == n - 1 + NST: into the slot just freed
          }
          mbar_wait_sleep(&v_full[s], (n / NST) & 1);
          const uint32_t va = smem_u32(sV + s * TILE_B);
          for (int g = 0; g < U.ntile; ++g) {
            if (j == 0 && items_g[g] > 0) mbar_wait_sleep(&o_free[g], (items_g[g] - 1) & 1);     // the previous item's O~ has been read
            mbar_wait_sleep(&p_full[g], pfull_n[g] & 1);
            ++pfull_n[g];
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < TK / 16; ++k)
              umma_bf16_ts(tmem_base + g * 256 + 192, tmem_base + g * 256 + 128 + k * 8, umma_desc_k128(va + k * 2048), idesc_o, (k != 0) || (j > 0));
            umma_commit(&o_full[g]);
          }
          umma_commit(&v_free[s]);
        }
        for (int g = 0; g < U.ntile; ++g) ++items_g[g];
        U = qt_item(p, U.w + stride);
      }
    }
    __syncwarp();
  } else {
    const int g = warp >> 2, q = warp & 3, row = q * 32 + lane;
    const uint32_t tS = tmem_base + ((uint32_t)(q * 32) << 16) + g * 256, tP = tS + 128, tO = tS + 192;
    const float sc = 0.125f * 1.4426950408889634f;              // 1/sqrt(64) * log2(e): scores in the log2 domain
    uint32_t nb = 0;                                             // running count of this tile's key blocks (barrier phases)
    for (QtItem U = qt_item(p, blockIdx.x); U.valid; U = qt_item(p, U.w + stride)) {
      if (g >= U.ntile) continue;
      const int Lk = U.Lk;
      float m = -INFINITY, l = 0.f;
      for (int j = 0; j < U.nkb; ++j, ++nb) {
        mbar_wait_sleep(&s_full[g], nb & 1);
        tc_fence_after();
        const int kbase = j * TK;
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld32_issue(tS + c * 32, s[c]);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[g]);
        if (kbase + TK > Lk) {       // keys past the row's length: -inf (exp2 -> 0)
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int k = 0; k < 32; ++k)
              if (kbase + c * 32 + k >= Lk) s[c][k] = 0xff800000u;
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
          for (int k = 0; k < 32; ++k) mx4[k & 3] = fmaxf(mx4[k & 3], __uint_as_float(s[c][k]));
        const float mb = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * sc;
        bool waited = j == 0;        // every phase of o_full is observed, in order (see the one-pass loop above)
        auto wait_prev = [&]() {
          if (!waited) { mbar_wait_sleep(&o_full[g], (nb - 1) & 1); tc_fence_after(); waited = true; }
        };
        const bool need = mb > m + 8.0f;
        if (__any_sync(0xffffffffu, need)) {
          const float m_new = need ? mb : m;
          if (j > 0) {
            const float f = ex2f(m - m_new);
            wait_prev();
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t r[16];
              tmem_ld16_issue(tO + c * 16, r);
              tmem_ld_wait();
#pragma unroll
              for (int k = 0; k < 16; ++k) r[k] = __float_as_uint(__uint_as_float(r[k]) * f);
              tmem_st16(tO + c * 16, r);
            }
            l *= f;
          }
          m = m_new;
        }
        float ls4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t pw[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float p0 = ex2f(fmaf(__uint_as_float(s[c][2 * e]), sc, -m)), p1 = ex2f(fmaf(__uint_as_float(s[c][2 * e + 1]), sc, -m));
            pw[e] = pack_bf16x2(p0, p1);
            ls4[e & 3] += __uint_as_float(pw[e] << 16) + __uint_as_float(pw[e] & 0xffff0000u);      // sum of the ROUNDED values
          }
          if (c == 0) wait_prev();
          tmem_st16(tP + c * 16, pw);
        }
        l += (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[g]);
      }
      // the item's last product: read O~ out, hand the columns back, then normalise and store from registers
      mbar_wait_sleep(&o_full[g], (nb - 1) & 1);
      tc_fence_after();
      float o[TD];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld32(tO + c * 32, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) o[c * 32 + k] = v[k];
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_free[g]);
      const float inv = 1.0f / l;
      const int t = U.q0 + g * TQ + row;
      if (t < p.L) {
        __nv_bfloat16* dst = p.out + ((size_t)U.b * p.L + t) * 512 + U.head * TD;
#pragma unroll
        for (int d8 = 0; d8 < TD / 8; ++d8)
          *reinterpret_cast<uint4*>(dst + d8 * 8) =
              make_uint4(pack_bf16x2(o[d8 * 8] * inv, o[d8 * 8 + 1] * inv), pack_bf16x2(o[d8 * 8 + 2] * inv, o[d8 * 8 + 3] * inv),
                         pack_bf16x2(o[d8 * 8 + 4] * inv, o[d8 * 8 + 5] * inv), pack_bf16x2(o[d8 * 8 + 6] * inv, o[d8 * 8 + 7] * inv));
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

static unsigned long long* g_atc_dbg = nullptr;
void attn_tc_set_debug(unsigned long long* buf) { g_atc_dbg = buf; }

template <int NG>
static int attn_tc_launch_as(const AttnTcParams& p, dim3 grid, cudaStream_t st) {
  static CbxPerDeviceOnce once;
  if (once.first()) {
    CBX_CUDA(cudaFuncSetAttribute(s3_attn_tc_kernel<NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, AtcCfg<NG>::SMEM));
    if (getenv("CBX_ATTN_OCC")) {      // debug aid: resident CTAs per SM as the runtime computes it
      int nb = 0;
      CBX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, s3_attn_tc_kernel<NG>, AtcCfg<NG>::THREADS, AtcCfg<NG>::SMEM));
      fprintf(stderr, "[cbx] s3_attn_tc_kernel<%d>: %d CTAs per SM (%d threads, %d B shared memory)\n", NG, nb, AtcCfg<NG>::THREADS, AtcCfg<NG>::SMEM);
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = dim3(AtcCfg<NG>::THREADS); cfg.dynamicSmemBytes = AtcCfg<NG>::SMEM; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_attn_tc_kernel<NG>, p));
  return CBX_OK;
}

int s3_attn_tc_launch(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st) {
  CBX_CHECK(ld_qkv == 1536, "s3_attn_tc: q|k|v row stride %d (expected 1536)", ld_qkv);
  AttnTcParams p;
  if (make_tmap_bf16_3d(&p.tm, qkv, 1536, L, B, 1536, (uint64_t)1536 * L, TD, TQ)) return CBX_ERR;
  p.out = out; p.lens = lens; p.L = L; p.nx = 0; p.n_items = 0;
  p.dbg = g_atc_dbg;
  if (g_atc_dbg) g_atc_dbg += 32;
  const dim3 grid((L + TQ - 1) / TQ, 8, B);
  // few CTAs (one utterance): two softmax groups per CTA split the key blocks; more CTAs than two per SM (padded batches): one
  // group per CTA and two CTAs per SM.  CBX_ATTN_GROUPS=1|2 forces a form.
  static const int force = [] { const char* v = getenv("CBX_ATTN_GROUPS"); return v ? atoi(v) : 0; }();
  const long long ctas = (long long)grid.x * grid.y * grid.z;
  const int ng = force == 1 || force == 2 ? force : (ctas > 2 * 148 ? 1 : 2);
  static const bool qt_on = [] { const char* v = getenv("CBX_ATTN_QT"); return !(v && v[0] == '0'); }();
  if (ng == 1 && qt_on && force == 0) {        // padded batches: two query tiles per CTA
    static CbxPerDeviceOnce once;
    if (once.first()) CBX_CUDA(cudaFuncSetAttribute(s3_attn_qt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, QtCfg::SMEM));
    cudaLaunchConfig_t cfg = {};
    p.nx = (int)(grid.x + 1) / 2; p.n_items = B * 8 * p.nx;
    cfg.gridDim = dim3(p.n_items < 148 ? p.n_items : 148); cfg.blockDim = dim3(QtCfg::THREADS); cfg.dynamicSmemBytes = QtCfg::SMEM; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_attn_qt_kernel, p));
    return CBX_OK;
  }
  return ng == 1 ? attn_tc_launch_as<1>(p, grid, st) : attn_tc_launch_as<2>(p, grid, st);
}

}  // namespace cbx
