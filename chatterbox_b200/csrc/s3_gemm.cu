// tcgen05 / TMEM / TMA GEMM with fused epilogues for the S3Gen flow-matching estimator
// (reference: src/chatterbox/models/s3gen/decoder.py:233-317 and matcha/transformer.py:243-316 —
// every Linear, the k=3 causal convolutions (as 3 row-shifted K panels accumulated in TMEM) and the
// 1x1 convolutions of ConditionalDecoder).
//
// One CTA computes a 128-row x n_tile (<= 256) output tile:
//   warp 0   : TMA producer — A tile [128 rows][64 k] and W tile [n_tile][64 k] per k-block, 128-byte
//              swizzle, 4-stage mbarrier ring; the causal taps are row-shifted loads of the same
//              activation tensor through a 3-D map [B][L][K], so rows t < 0 arrive as zeros;
//   warp 1   : MMA issuer — one elected thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32,
//              M = 128, N = n_tile, K = 16) x 4 per stage, tcgen05.commit frees the stage / signals
//              the epilogue; also owns the TMEM allocation;
//   warps 2-5: drain — each thread owns one accumulator row (TMEM lane) and copies it with tcgen05.ld to an
//              fp32 tile in shared memory (the operand ring, idle by then);
//   warps 6-15 stage the per-column epilogue vectors meanwhile; then all 16 warps run the epilogue with the
//              lanes across columns (coalesced): bias, LayerNorm-over-channels + Mish (+ time embedding),
//              GELU / ELU, residual adds, fp32 / bf16 stores (Snake / leaky-ReLU), and the LayerNorm
//              feeding the next GEMM.
#include "s3_gemm.cuh"
#include <cstdlib>
#include "../../include/cbx.h"

namespace cbx {

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

int make_tmap_bf16_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t ld1,
                      uint64_t ld2, uint32_t b0, uint32_t b1) {
  PFN_encodeTiled enc = get_encode_tiled();
  CBX_CHECK(enc, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {ld1 * 2, ld2 * 2};
  cuuint32_t box[3] = {b0, b1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CBX_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) dims %llu %llu %llu", (int)r,
            (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2);
  return CBX_OK;
}

namespace {

constexpr int BM = 128, BK = 64, STAGES = 4;         // STAGES: ring depth at the widest tile (n_tile = 256)
constexpr int MAX_STAGES = 8;                        // narrower tiles get a deeper ring in the same shared memory
constexpr int A_BYTES = BM * BK * 2;                 // 16 KB
constexpr int W_BYTES_MAX = 256 * BK * 2;            // 32 KB
constexpr int STAGE_BYTES = A_BYTES + W_BYTES_MAX;
constexpr int GEMM_THREADS = 512;
constexpr int MAX_SPLIT = 4;                         // largest split-K cluster
constexpr int EPI_R = 2;                             // rows per epilogue iteration and warp
enum { PAR_BIAS = 0, PAR_G1, PAR_B1, PAR_TEMB, PAR_G2, PAR_B2, PAR_ALPHA, PAR_COUNT };   // per-column epilogue vectors
constexpr int PAR_BYTES = PAR_COUNT * 256 * 4;
constexpr int STAT_BYTES = MAX_SPLIT * BM * 2 * 4;   // N-split LayerNorm: per-row (sum, sum of squares) pushed by every CTA of the cluster
constexpr int GEMM_SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + PAR_BYTES + STAT_BYTES;

struct GemmParams {
  CUtensorMap tmA, tmW;
  int B, L, N, K, taps, n_tile, tiles_l, dil, pad, M_out, ksplit, hilo, nsplit;
  int nstages, stage_bytes;   // ring geometry: stage = A tile (16 KB) + W tile (n_tile x 128 B), 1024-byte aligned
  GemmEpi e;
  unsigned long long* dbg;    // optional %globaltimer stamps of CTA (0,0) (debug aid)
};

// mish(x) = x tanh(softplus(x)) = x t / (t + 2) with t = e^x (e^x + 2); one ex2 and one fast division
__device__ __forceinline__ float mish_f(float x) {
  const float n = __expf(fminf(x, 20.f));
  const float t = n * (n + 2.0f);
  return x * __fdividef(t, t + 2.0f);
}
// sin with a two-constant Cody-Waite reduction to [-pi, pi] ahead of the SFU approximation
__device__ __forceinline__ float sin_f(float x) {
  const float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(k, -6.28318548202514648f, x);
  r = fmaf(k, 1.74845553e-7f, r);
  return __sinf(r);
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define GEMM_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) p.dbg[k] = gtimer(); } while (0)

// FEAT: the epilogue features an instantiation carries code for.  The epilogue body runs cold once per CTA, so
// what it costs is instruction fetch over everything on its path — dispatching to the narrowest instantiation
// (gemm_launch) keeps the common estimator epilogues short.
enum : uint32_t { F_LN1 = 1, F_GELU = 2, F_ELU = 4, F_RESID = 8, F_RESID2 = 16, F_F32 = 32, F_BF16 = 64, F_LN2 = 128,
                  F_ACT = 256 /* snake / leaky-relu on the bf16 output */, F_SHIFT = 512, F_HILO = 1024, F_SWIGLU = 2048,
                  F_ALL = 4095 };

template <uint32_t FEAT>
__global__ void __launch_bounds__(GEMM_THREADS, 1) s3_gemm_kernel(const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* accum_bar = empty + MAX_STAGES;
  const int NS = p.nstages, SB = p.stage_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  uint64_t* stat_bar = accum_bar + 2;        // N-split: completes when every CTA's row moments have landed in our shared memory
  float* par = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES + 256);
  float2* stat = reinterpret_cast<float2*>(smem + STAGES * STAGE_BYTES + 256 + PAR_BYTES);
  float* tile = reinterpret_cast<float*>(smem);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();      // the next kernel may be scheduled: it waits (pdl_wait) for this grid to finish
  if (threadIdx.x == 0) GEMM_STAMP(0);
  const int b = blockIdx.x / p.tiles_l, t0 = (blockIdx.x % p.tiles_l) * BM;
  const int n0 = blockIdx.y * p.n_tile;
  // split-K over a cluster (1, 1, S): CTA `rank` accumulates K panels [kb0, kb0 + nkb) of the tile into its own
  // TMEM, the partial tiles meet in distributed shared memory and each CTA finishes 128 / S rows of the tile
  const int S = p.ksplit, rank = blockIdx.z;
  const int NSP = p.nsplit;                  // N-split cluster (1, NSP, 1): LayerNorm moments are exchanged, S == 1
  // hi/lo mode (T3): a k-block is ONE weight panel and BOTH activation planes (hi, lo) multiplied against it
  const int kpb = p.K / BK, nkb_all = p.hilo ? kpb : p.taps * kpb;
  const uint32_t w_bytes = (uint32_t)p.n_tile * BK * 2;
  const int nkb = nkb_all / S + (rank < nkb_all % S ? 1 : 0);                 // >= 1 (host: nkb_all >= S)
  const int kb0 = rank * (nkb_all / S) + min(rank, nkb_all % S);
  const int row_base = rank * (BM / S), rpw = 8 / S;        // epilogue rows of this CTA / per warp
  const int pitch = p.n_tile + 4;            // floats: row-per-lane float4 stores stay bank-conflict free
  const uint32_t tmem_cols = p.n_tile > 128 ? 256u : (p.n_tile > 64 ? 128u : (p.n_tile > 32 ? 64u : 32u));

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum_bar, 1);
    mbar_init(stat_bar, 1);
    // N-split: the cluster's CTAs deliver BM rows x (sum, sum of squares) each with st.async, completing bytes on this barrier
    if (p.nsplit > 1) mbar_arrive_expect_tx(stat_bar, (uint32_t)(p.nsplit * BM * 8));
    mbar_fence_init();
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmW);
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // N-split: peers push into our shared memory / arrive on stat_bar later; they may only do so once it is initialised.
  // Arrive now, wait just before the first remote access (the barrier completes somewhere behind the main loop).
  if (p.nsplit > 1) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  if (threadIdx.x == 0) GEMM_STAMP(1);
  // PDL: everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel.  Warp 0 goes on to
  // issue the WEIGHT panels (constant data) before it waits; every other warp waits here, before its first read of
  // anything an earlier kernel wrote (residual rows, time-embedding vector).
  if (warp != 0) pdl_wait();

  // epilogue state shared by all warps (see phase 2 below)
  const GemmEpi& e = p.e;
  const bool has_ln1 = (FEAT & F_LN1) && e.ln_g != nullptr, has_gelu = (FEAT & F_GELU) && e.gelu != 0;
  const bool has_elu = (FEAT & F_ELU) && e.elu != 0, has_resid = (FEAT & F_RESID) && e.resid != nullptr;
  const bool has_resid2 = (FEAT & F_RESID2) && e.resid2 != nullptr;
  const bool has_f32 = (FEAT & F_F32) && e.out_f32 != nullptr, has_bf16 = (FEAT & F_BF16) && e.out_bf16 != nullptr;
  const bool has_ln2 = (FEAT & F_LN2) && e.out_ln != nullptr, has_snake = (FEAT & F_ACT) && e.snake_alpha != nullptr;
  const bool has_hilo = (FEAT & F_HILO) && e.out_hilo != nullptr, has_swiglu = (FEAT & F_SWIGLU) && e.swiglu != 0;
  const float lrelu = (FEAT & F_ACT) ? e.lrelu : 0.f;
  const long long out_shift = (FEAT & F_SHIFT) ? e.out_shift : 0, out_limit = (FEAT & F_SHIFT) ? e.out_limit : 0;
  const int ngrp = p.n_tile / 4;
  const bool g_ok[2] = {lane < ngrp, lane + 32 < ngrp};
  const int gcol[2] = {lane * 4, g_ok[1] ? (lane + 32) * 4 : 0};
  float4 rvn[EPI_R][2];
  auto load_resid = [&](int r0) {
#pragma unroll
    for (int rr = 0; rr < EPI_R; ++rr) {
      const int t = t0 + r0 + rr;
      const size_t gr = (size_t)b * p.M_out + (t < p.M_out ? t : 0);
#pragma unroll
      for (int i = 0; i < 2; ++i)
        rvn[rr][i] = (t < p.M_out && g_ok[i]) ? *reinterpret_cast<const float4*>(e.resid + gr * e.ld_resid + n0 + gcol[i])
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
#pragma unroll
  for (int rr = 0; rr < EPI_R; ++rr) { rvn[rr][0] = make_float4(0.f, 0.f, 0.f, 0.f); rvn[rr][1] = rvn[rr][0]; }
  if (has_resid && warp != 0 && NSP == 1) load_resid(row_base + warp * rpw);
  // N-split ownership: thread <-> (row 8 warp + lane / 4, 16 columns): its residual values are requested now
  float4 nres[4];
  const int ns_r = warp * 8 + (lane >> 2), ns_c0 = (lane & 3) * 16;
  const bool ns_ok = t0 + ns_r < p.M_out;
  const size_t ns_gr = (size_t)b * p.M_out + (ns_ok ? t0 + ns_r : 0);
  auto ns_prefetch = [&]() {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      nres[i] = (has_resid && ns_ok) ? *reinterpret_cast<const float4*>(e.resid + ns_gr * e.ld_resid + n0 + ns_c0 + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  if (NSP > 1 && warp != 0) ns_prefetch();

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t tx = (p.hilo ? 2 : 1) * A_BYTES + w_bytes;
      const int npre = nkb < NS ? nkb : NS;
      // weight panels of the first ring pass: issued before the PDL wait (they do not depend on the previous kernel)
      for (int i = 0; i < npre; ++i) {
        const int kb = kb0 + i;
        mbar_arrive_expect_tx(&full[i], tx);
        tma_load_3d(smem + i * SB + A_BYTES, &p.tmW, &full[i], kb * BK, n0, 0);
      }
      pdl_wait();
      GEMM_STAMP(11);
      // ring slot / phase and (tap, k-block-in-tap) advance as counters: this loop is one thread's serial code, a runtime
      // integer division per index costs ~0.1 us per k-block on the critical path
      int s = 0, ph = 0, tap = kb0 / kpb, kk = kb0 % kpb;
      for (int i = 0; i < nkb; ++i) {
        const int kb = kb0 + i;
        uint8_t* sa = smem + s * SB;
        if (i >= npre) {
          mbar_wait_spin(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], tx);
          tma_load_3d(sa + A_BYTES, &p.tmW, &full[s], kb * BK, n0, 0);
        }
        if (p.hilo) {              // planes 0 (hi) and 1 (lo) of A; the lo tile sits behind the W tile
          tma_load_3d(sa, &p.tmA, &full[s], kb * BK, t0, 0);
          tma_load_3d(sa + A_BYTES + w_bytes, &p.tmA, &full[s], kb * BK, t0, 1);
        } else tma_load_3d(sa, &p.tmA, &full[s], kk * BK, t0 + tap * p.dil - p.pad, b);
        if (++s == NS) { s = 0; ph ^= 1; }
        if (++kk == kpb) { kk = 0; ++tap; }
      }
      GEMM_STAMP(2);
    }
    if (NSP > 1) { __syncwarp(); pdl_wait(); ns_prefetch(); }      // the producer warp's own residual rows, once its loads are out
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(BM, p.n_tile);
      int s = 0, ph = 0;
      for (int kb = 0; kb < nkb; ++kb) {      // kb: local panel index
        mbar_wait_spin(&full[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * SB);
        const uint64_t da = umma_desc_k128(sa), db = umma_desc_k128(sa + A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)    // +32 bytes per K=16 step inside the 128-byte swizzle row
          umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
        if (p.hilo) {
          const uint64_t da1 = umma_desc_k128(sa + A_BYTES + w_bytes);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) umma_bf16(tmem_base, da1 + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, 1u);
        }
        umma_commit(&empty[s]);
        if (++s == NS) { s = 0; ph ^= 1; }
      }
      umma_commit(accum_bar);
      GEMM_STAMP(3);
    }
  } else if (warp >= 6) {
    // idle until the epilogue: stage the per-column epilogue vectors of this tile in shared memory
    for (int c = threadIdx.x - 6 * 32; c < p.n_tile; c += GEMM_THREADS - 6 * 32) {
      const int n = n0 + c;
      par[PAR_BIAS * 256 + c] = e.bias ? __ldg(e.bias + (e.bias_mod > 0 ? n % e.bias_mod : n)) : 0.f;
      par[PAR_G1 * 256 + c] = e.ln_g ? __ldg(e.ln_g + n) : 0.f;
      par[PAR_B1 * 256 + c] = e.ln_g ? __ldg(e.ln_b + n) : 0.f;
      par[PAR_TEMB * 256 + c] = e.temb ? __ldg(e.temb + n) : 0.f;      // one vector [N]: every row of a solve shares t
      par[PAR_G2 * 256 + c] = e.out_ln ? __ldg(e.ln2_g + n) : 0.f;
      par[PAR_B2 * 256 + c] = e.out_ln ? __ldg(e.ln2_b + n) : 0.f;
      par[PAR_ALPHA * 256 + c] = e.snake_alpha ? __ldg(e.snake_alpha + n) : 1.f;
    }
  }
  __syncwarp();                                   // warps 0 / 1: lane 0 rejoins
  {
    // ===== epilogue, phase 1 (ALL 16 warps): drain the accumulator, thread <-> row (TMEM lane), to an fp32 tile in
    // shared memory (the operand ring is idle once accum_bar fires).  A warp may touch the 32 TMEM lanes of its
    // quarter (warp % 4) only, so the four warps of a quarter take every fourth 16-column chunk each. =====
    const int q = warp & 3, row = q * 32 + lane;
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
    // GEMM_DRAIN: 0 = warps 2-5 only (round 1); 1 = all warps, each spinning on the accumulator barrier; 2 (default) = all
    // warps, but only warps 2-5 wait on the mbarrier while the others sleep in a CTA barrier (no issue slots taken from
    // the producer / MMA lanes during the main loop)
#ifndef GEMM_DRAIN
#define GEMM_DRAIN 2
#endif
#if GEMM_DRAIN == 2
    if (warp >= 2 && warp <= 5) mbar_wait_spin(accum_bar, 0);
    __syncthreads();
#elif GEMM_DRAIN == 1
    mbar_wait_spin(accum_bar, 0);
#else
    if (warp >= 2 && warp <= 5) mbar_wait_spin(accum_bar, 0);
#endif
    tc_fence_after();
    if (threadIdx.x == 64) GEMM_STAMP(4);
    const int nch = p.n_tile / 16;
#if GEMM_DRAIN == 0
    if (warp >= 2 && warp <= 5)
    for (int c = 0; c < nch; ++c) {
#else
    for (int c = warp >> 2; c < nch; c += 4) {
#endif
      float v[16];
      tmem_ld16(trow + c * 16, v);
      float4* dst = reinterpret_cast<float4*>(tile + (size_t)row * pitch + c * 16);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    tc_fence_before();
    if (threadIdx.x == 64) GEMM_STAMP(5);
  }
  __syncwarp();
  if (S > 1) cluster_sync_all(); else __syncthreads();
  if (threadIdx.x == 64) GEMM_STAMP(8);
  const float inv_n = 1.0f / (float)(p.n_tile * NSP);
  // ===== epilogue, phase 2 (all 16 warps): warp <-> 8 rows, lanes across float4 column groups (lane, lane + 32),
  // two rows per iteration of a ROLLED loop (the body is executed cold once per CTA, so instruction fetch — not
  // issue — is what a large unrolled body costs); the residual rows of the next iteration are loaded one
  // iteration ahead, the first pair before the barrier (behind the main loop).  Every mode test is on a kernel
  // parameter, hoisted out of the element loops. =====
  if (NSP > 1) {
    // ===== N-split epilogue (cluster (1, NSP, 1), n_tile = 64): the LayerNorm of a row spans the NSP column tiles of the
    // cluster.  A thread keeps its 16 values of the tile in registers (row = 8 warp + lane / 4, columns 16 (lane % 4) ...):
    // partial (sum, sum of squares) over this CTA's 64 columns are PUSHED into every CTA of the cluster (st.async: the store
    // completes bytes on the receiver's mbarrier); a CTA waits on its own mbarrier, adds the NSP partial moments from its own shared
    // memory, normalises and stores.  Nothing is read twice, no cluster barrier on the path, no trailing barrier before exit
    // (a peer writes here only before its arrive, and we leave only after all arrives). =====
    if constexpr ((FEAT & (F_LN1 | F_LN2)) != 0) {
      const int r = ns_r, c0 = ns_c0;
      const bool ok = ns_ok;
      const size_t gr = ns_gr;
      float v[16];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 a = *reinterpret_cast<const float4*>(tile + (size_t)r * pitch + c0 + 4 * i);
        const float4 bb = *reinterpret_cast<const float4*>(par + PAR_BIAS * 256 + c0 + 4 * i);
        v[4 * i] = a.x + bb.x; v[4 * i + 1] = a.y + bb.y; v[4 * i + 2] = a.z + bb.z; v[4 * i + 3] = a.w + bb.w;
      }
      auto add_res = [&]() {
#pragma unroll
        for (int i = 0; i < 4; ++i) { v[4 * i] += nres[i].x; v[4 * i + 1] += nres[i].y; v[4 * i + 2] += nres[i].z; v[4 * i + 3] += nres[i].w; }
      };
      auto store_out = [&]() {
        if (has_f32 && ok) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<float4*>(e.out_f32 + gr * e.ld_f32 + n0 + c0 + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        if (has_bf16 && ok) {
#pragma unroll
          for (int i = 0; i < 2; ++i)
            *reinterpret_cast<uint4*>(e.out_bf16 + gr * e.ld_bf16 + n0 + c0 + 8 * i) =
                make_uint4(pack_bf16x2(v[8 * i], v[8 * i + 1]), pack_bf16x2(v[8 * i + 2], v[8 * i + 3]),
                           pack_bf16x2(v[8 * i + 4], v[8 * i + 5]), pack_bf16x2(v[8 * i + 6], v[8 * i + 7]));
        }
      };
      if (!has_ln1) {          // LayerNorm of the final value: residual adds first
        if (has_resid) add_res();
        if (has_resid2 && ok) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 a = *reinterpret_cast<const float4*>(e.resid2 + gr * e.ld_resid2 + n0 + c0 + 4 * i);
            v[4 * i] += a.x; v[4 * i + 1] += a.y; v[4 * i + 2] += a.z; v[4 * i + 3] += a.w;
          }
        }
      }
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int k = 0; k < 16; ++k) { s1 += v[k]; s2 += v[k] * v[k]; }
      s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s2 += __shfl_xor_sync(0xffffffffu, s2, 1);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 2); s2 += __shfl_xor_sync(0xffffffffu, s2, 2);
      if (threadIdx.x == 64) GEMM_STAMP(9);
      // push the row's partial moments into every CTA of the cluster (slot = our rank), then one arrive per warp and peer
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");        // every peer's stat_bar is initialised
      const int my = blockIdx.y % NSP;
      if ((lane & 3) == 0) {
        // st.async: the 8 bytes land in the peer's shared memory and complete on ITS mbarrier — data and signal in one
        // fire-and-forget operation, no release fence / remote arrive round trips on this thread
        const uint32_t la = smem_u32(&stat[my * BM + r]), lb = smem_u32(stat_bar);
        for (int j = 0; j < NSP; ++j)
          asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                       ::"r"(dsmem_addr(la, (uint32_t)j)), "f"(s1), "f"(s2), "r"(dsmem_addr(lb, (uint32_t)j)) : "memory");
      }
      if (threadIdx.x == 64) GEMM_STAMP(10);
      if (!has_ln1) store_out();          // the fp32 / bf16 copies of the un-normalised value leave while the moments travel
      if (threadIdx.x == 64) GEMM_STAMP(12);
      {
        unsigned spins = 0;
        uint32_t done = 0;
        while (!done) {
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(done) : "r"(smem_u32(stat_bar)), "r"(0u) : "memory");
          if (++spins > (1u << 26)) __trap();
        }
      }
      if (threadIdx.x == 64) GEMM_STAMP(13);
      s1 = 0.f; s2 = 0.f;
#pragma unroll
      for (int j = 0; j < MAX_SPLIT; ++j)
        if (j < NSP) { const float2 m2 = stat[j * BM + r]; s1 += m2.x; s2 += m2.y; }
      const float mean = s1 * inv_n;
      const float rstd = rsqrtf(fmaxf(s2 * inv_n - mean * mean, 0.f) + 1e-5f);
      if (has_ln1) {           // LayerNorm + Mish (+ time embedding) -> bf16 and / or fp32
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 g4 = *reinterpret_cast<const float4*>(par + PAR_G1 * 256 + c0 + 4 * i);
          const float4 b4 = *reinterpret_cast<const float4*>(par + PAR_B1 * 256 + c0 + 4 * i);
          const float4 t4 = *reinterpret_cast<const float4*>(par + PAR_TEMB * 256 + c0 + 4 * i);
          v[4 * i] = mish_f((v[4 * i] - mean) * rstd * g4.x + b4.x) + t4.x;
          v[4 * i + 1] = mish_f((v[4 * i + 1] - mean) * rstd * g4.y + b4.y) + t4.y;
          v[4 * i + 2] = mish_f((v[4 * i + 2] - mean) * rstd * g4.z + b4.z) + t4.z;
          v[4 * i + 3] = mish_f((v[4 * i + 3] - mean) * rstd * g4.w + b4.w) + t4.w;
        }
        if (has_resid) add_res();
        store_out();
      } else if (has_ln2 && ok) {
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 g4 = *reinterpret_cast<const float4*>(par + PAR_G2 * 256 + c0 + 4 * i);
          const float4 b4 = *reinterpret_cast<const float4*>(par + PAR_B2 * 256 + c0 + 4 * i);
          w[2 * i] = pack_bf16x2((v[4 * i] - mean) * rstd * g4.x + b4.x, (v[4 * i + 1] - mean) * rstd * g4.y + b4.y);
          w[2 * i + 1] = pack_bf16x2((v[4 * i + 2] - mean) * rstd * g4.z + b4.z, (v[4 * i + 3] - mean) * rstd * g4.w + b4.w);
        }
        *reinterpret_cast<uint4*>(e.out_ln + gr * e.ld_ln + n0 + c0) = make_uint4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<uint4*>(e.out_ln + gr * e.ld_ln + n0 + c0 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
      }
    }
  } else {
    if (warp == 0) {            // the producer warp skipped the early residual prefetch (it was issuing TMA loads)
      pdl_wait();
      if (has_resid) load_resid(row_base);
    }
    float4 bias4[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) bias4[i] = g_ok[i] ? *reinterpret_cast<const float4*>(par + PAR_BIAS * 256 + gcol[i]) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int it = 0; it < rpw / EPI_R; ++it) {
      const int r0 = row_base + warp * rpw + it * EPI_R;
      float x[EPI_R][2][4];
      bool okr[EPI_R];
      size_t grow[EPI_R];
      float4 rv[EPI_R][2];
#pragma unroll
      for (int rr = 0; rr < EPI_R; ++rr) {
        const int t = t0 + r0 + rr;
        okr[rr] = t < p.M_out;
        grow[rr] = (size_t)b * p.M_out + (okr[rr] ? t : 0);
#pragma unroll
        for (int i = 0; i < 2; ++i) rv[rr][i] = rvn[rr][i];
      }
      if (has_resid && it + 1 < rpw / EPI_R) load_resid(r0 + EPI_R);
      if (S == 1) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const float4 v = g_ok[i] ? *reinterpret_cast<const float4*>(tile + (size_t)(r0 + rr) * pitch + gcol[i]) : make_float4(0.f, 0.f, 0.f, 0.f);
            x[rr][i][0] = v.x + bias4[i].x; x[rr][i][1] = v.y + bias4[i].y; x[rr][i][2] = v.z + bias4[i].z; x[rr][i][3] = v.w + bias4[i].w;
          }
      } else {
        // partial tiles of the S cluster CTAs: all loads of a row pair issued before the first add; summed in rank
        // order, so the result does not depend on which CTA finishes the row
        float4 u[EPI_R][2][MAX_SPLIT];
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const uint32_t la = smem_u32(tile + (size_t)(r0 + rr) * pitch + gcol[i]);
#pragma unroll
            for (int j = 0; j < MAX_SPLIT; ++j)
              u[rr][i][j] = (g_ok[i] && j < S) ? dsmem_ld4(dsmem_addr(la, (uint32_t)j)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            float4 v = u[rr][i][0];
#pragma unroll
            for (int j = 1; j < MAX_SPLIT; ++j) { v.x += u[rr][i][j].x; v.y += u[rr][i][j].y; v.z += u[rr][i][j].z; v.w += u[rr][i][j].w; }
            x[rr][i][0] = v.x + bias4[i].x; x[rr][i][1] = v.y + bias4[i].y; x[rr][i][2] = v.z + bias4[i].z; x[rr][i][3] = v.w + bias4[i].w;
          }
      }
      if (threadIdx.x == 64 && it == 0) { if (x[0][0][0] == 123.456f) GEMM_STAMP(15); GEMM_STAMP(9); }
      if (has_ln1) {
        float s1[EPI_R], s2[EPI_R];
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr) {
          s1[rr] = 0.f; s2[rr] = 0.f;
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) { s1[rr] += x[rr][i][k]; s2[rr] += x[rr][i][k] * x[rr][i][k]; }   // masked groups hold zeros
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
          for (int rr = 0; rr < EPI_R; ++rr) {
            s1[rr] += __shfl_xor_sync(0xffffffffu, s1[rr], o);
            s2[rr] += __shfl_xor_sync(0xffffffffu, s2[rr], o);
          }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (!g_ok[i]) continue;
          const float4 g = *reinterpret_cast<const float4*>(par + PAR_G1 * 256 + gcol[i]);
          const float4 bb = *reinterpret_cast<const float4*>(par + PAR_B1 * 256 + gcol[i]);
          const float4 te = *reinterpret_cast<const float4*>(par + PAR_TEMB * 256 + gcol[i]);
          const float gg[4] = {g.x, g.y, g.z, g.w}, bv[4] = {bb.x, bb.y, bb.z, bb.w}, tv[4] = {te.x, te.y, te.z, te.w};
#pragma unroll
          for (int rr = 0; rr < EPI_R; ++rr) {
            const float mean = s1[rr] * inv_n;
            const float rstd = rsqrtf(fmaxf(s2[rr] * inv_n - mean * mean, 0.f) + 1e-5f);
#pragma unroll
            for (int k = 0; k < 4; ++k) x[rr][i][k] = mish_f((x[rr][i][k] - mean) * rstd * gg[k] + bv[k]) + tv[k];
          }
        }
      }
      if (has_gelu) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) x[rr][i][k] = gelu_f(x[rr][i][k]);
      }
      if (has_elu) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) x[rr][i][k] = x[rr][i][k] > 0.f ? x[rr][i][k] : __expf(x[rr][i][k]) - 1.0f;
      }
      if (has_resid) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) { x[rr][i][0] += rv[rr][i].x; x[rr][i][1] += rv[rr][i].y; x[rr][i][2] += rv[rr][i].z; x[rr][i][3] += rv[rr][i].w; }
      }
      if (has_resid2) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i)
            rv[rr][i] = (okr[rr] && g_ok[i]) ? *reinterpret_cast<const float4*>(e.resid2 + grow[rr] * e.ld_resid2 + n0 + gcol[i])
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) { x[rr][i][0] += rv[rr][i].x; x[rr][i][1] += rv[rr][i].y; x[rr][i][2] += rv[rr][i].z; x[rr][i][3] += rv[rr][i].w; }
      }
      if (threadIdx.x == 64 && it == 0) { if (x[0][0][0] == 123.456f) GEMM_STAMP(15); GEMM_STAMP(10); }
      if (has_swiglu) {    // group 0 = gate columns, group 1 = the matching up columns (n_tile == 256)
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float gte = x[rr][0][k];
            x[rr][0][k] = __fdividef(gte, 1.0f + __expf(-gte)) * x[rr][1][k];
          }
      }
      if (has_hilo) {
        const int ngroups = has_swiglu ? 1 : 2;
        const int ncol0 = has_swiglu ? (n0 >> 1) : n0;
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            if (!okr[rr] || !g_ok[i] || i >= ngroups) continue;
            uint32_t h0, l0, h1, l1;
            split_hi_lo(x[rr][i][0], x[rr][i][1], h0, l0);
            split_hi_lo(x[rr][i][2], x[rr][i][3], h1, l1);
            __nv_bfloat16* dsth = e.out_hilo + grow[rr] * e.ld_hilo + ncol0 + gcol[i];
            *reinterpret_cast<uint2*>(dsth) = make_uint2(h0, h1);
            *reinterpret_cast<uint2*>(dsth + e.hilo_plane) = make_uint2(l0, l1);
          }
      }
      if (has_f32) {
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            if (!okr[rr] || !g_ok[i]) continue;
            const long long off = (long long)(grow[rr] * e.ld_f32 + n0 + gcol[i]) + out_shift;
            if (out_limit == 0 || (off >= 0 && off + 4 <= out_limit))
              *reinterpret_cast<float4*>(e.out_f32 + off) = make_float4(x[rr][i][0], x[rr][i][1], x[rr][i][2], x[rr][i][3]);
          }
      }
      if (has_bf16) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (!g_ok[i]) continue;
          float al[4] = {1.f, 1.f, 1.f, 1.f}, ial[4] = {1.f, 1.f, 1.f, 1.f};
          if (has_snake) {
            const float4 a4 = *reinterpret_cast<const float4*>(par + PAR_ALPHA * 256 + gcol[i]);
            al[0] = a4.x; al[1] = a4.y; al[2] = a4.z; al[3] = a4.w;
#pragma unroll
            for (int k = 0; k < 4; ++k) ial[k] = 1.0f / (al[k] + 1e-9f);
          }
#pragma unroll
          for (int rr = 0; rr < EPI_R; ++rr) {
            if (!okr[rr]) continue;
            float y[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              y[k] = x[rr][i][k];
              if (has_snake) { const float sn = sin_f(y[k] * al[k]); y[k] += ial[k] * sn * sn; }
              else if (lrelu != 0.f) y[k] = y[k] > 0.f ? y[k] : y[k] * lrelu;
            }
            *reinterpret_cast<uint2*>(e.out_bf16 + grow[rr] * e.ld_bf16 + n0 + gcol[i]) = make_uint2(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]));
          }
        }
      }
      if (has_ln2) {
        float s1[EPI_R], s2[EPI_R];
#pragma unroll
        for (int rr = 0; rr < EPI_R; ++rr) {
          s1[rr] = 0.f; s2[rr] = 0.f;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            if (!g_ok[i]) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k) { s1[rr] += x[rr][i][k]; s2[rr] += x[rr][i][k] * x[rr][i][k]; }
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
          for (int rr = 0; rr < EPI_R; ++rr) {
            s1[rr] += __shfl_xor_sync(0xffffffffu, s1[rr], o);
            s2[rr] += __shfl_xor_sync(0xffffffffu, s2[rr], o);
          }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (!g_ok[i]) continue;
          const float4 g = *reinterpret_cast<const float4*>(par + PAR_G2 * 256 + gcol[i]);
          const float4 bb = *reinterpret_cast<const float4*>(par + PAR_B2 * 256 + gcol[i]);
#pragma unroll
          for (int rr = 0; rr < EPI_R; ++rr) {
            if (!okr[rr]) continue;
            const float mean = s1[rr] * inv_n;
            const float rstd = rsqrtf(fmaxf(s2[rr] * inv_n - mean * mean, 0.f) + 1e-5f);
            const float y0 = (x[rr][i][0] - mean) * rstd * g.x + bb.x, y1 = (x[rr][i][1] - mean) * rstd * g.y + bb.y;
            const float y2 = (x[rr][i][2] - mean) * rstd * g.z + bb.z, y3 = (x[rr][i][3] - mean) * rstd * g.w + bb.w;
            *reinterpret_cast<uint2*>(e.out_ln + grow[rr] * e.ld_ln + n0 + gcol[i]) = make_uint2(pack_bf16x2(y0, y1), pack_bf16x2(y2, y3));
          }
        }
      }
    }
  }
  if (threadIdx.x == 64) GEMM_STAMP(6);
  __syncwarp();
  if (S > 1) cluster_sync_all(); else __syncthreads();     // split-K: peers may still be reading this CTA's partial tile
  if (threadIdx.x == 0) GEMM_STAMP(7);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

}  // namespace

int gemm_make(GemmDesc* d, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int B, int L, int N, int K, int taps,
              const GemmEpi& epi) {
  return gemm_make_conv(d, A, lda, W, B, L, N, K, taps, 1, taps - 1, L, epi);
}

int gemm_make_conv(GemmDesc* d, const __nv_bfloat16* A, int lda, const __nv_bfloat16* W, int B, int L, int N, int K, int taps,
                   int dil, int pad, int M_out, const GemmEpi& epi) {
  CBX_CHECK(K % BK == 0 && K >= BK, "gemm_make: K=%d must be a multiple of 64", K);
  int n_tile = 0;
  for (int c = 256; c >= 16; c -= 16) if (N % c == 0) { n_tile = c; break; }
  CBX_CHECK(n_tile > 0, "gemm_make: N=%d unsupported (needs a divisor that is a multiple of 16, <= 256)", N);
  CBX_CHECK(B == 1 || M_out == L, "gemm_make: M_out != L needs B == 1");
  // a row LayerNorm over N = 256 channels: four 64-column tiles in one cluster exchange the row moments (every SM ingests
  // a quarter of the weight panel and drains a quarter of the tile; no split-K partial tiles through DSMEM).
  // CBX_GEMM_NSPLIT=1 restores the full-row tile + split-K form.
  static const int nsplit_max = [] { const char* v = getenv("CBX_GEMM_NSPLIT"); return v ? atoi(v) : 4; }();
  d->nsplit = 1;
  const bool simple_epi = !epi.swiglu && !epi.out_hilo && !epi.gelu && !epi.elu && !epi.snake_alpha && epi.lrelu == 0.f &&
                          epi.out_shift == 0 && epi.out_limit == 0 && epi.bias_mod == 0;
  if ((epi.ln_g != nullptr) != (epi.out_ln != nullptr) && n_tile == N && N == 256 && nsplit_max == 4 && simple_epi &&
      B * ((M_out + BM - 1) / BM) * nsplit_max <= 296) {
    d->nsplit = nsplit_max; n_tile = N / nsplit_max;
  }
  CBX_CHECK(!(epi.ln_g || epi.out_ln) || n_tile * d->nsplit == N, "gemm_make: row LayerNorm needs the whole row in one tile / cluster (N=%d)", N);
    d->B = B; d->L = L; d->N = N; d->K = K; d->taps = taps; d->n_tile = n_tile; d->epi = epi;
  d->dil = dil; d->pad = pad; d->M_out = M_out; d->hilo = 0;
  // few output tiles (the estimator's 2 x ceil(L / 128) row tiles): spread the K panels of each tile over a cluster
  const int tiles = B * ((M_out + BM - 1) / BM) * (N / n_tile), nkb_all = taps * (K / BK);
  d->ksplit = 1;
  static const int max_split = [] { const char* v = getenv("CBX_GEMM_KSPLIT"); return v ? atoi(v) : 2; }();
  for (int sp = max_split; sp >= 2 && d->nsplit == 1; sp >>= 1)
    if (nkb_all >= sp && tiles * sp <= 148) { d->ksplit = sp; break; }
  if (make_tmap_bf16_3d(&d->tmA, A, K, L, B, lda, (uint64_t)lda * L, BK, BM)) return CBX_ERR;
  if (make_tmap_bf16_3d(&d->tmW, W, (uint64_t)taps * K, N, 1, (uint64_t)taps * K, (uint64_t)taps * K * N, BK, n_tile)) return CBX_ERR;
  return CBX_OK;
}

int gemm_make_hilo(GemmDesc* d, const __nv_bfloat16* A, int lda, long long plane, const __nv_bfloat16* W, int M, int N, int K,
                   const GemmEpi& epi) {
  return gemm_make_hilo_ex(d, A, lda, plane, W, M, N, N, K, epi, 0, 0);
}

int gemm_make_hilo_ex(GemmDesc* d, const __nv_bfloat16* A, int lda, long long plane, const __nv_bfloat16* W, int M, int N, int N_w,
                      int K, const GemmEpi& epi, int n_tile_req, int ksplit_req) {
  CBX_CHECK(K % BK == 0 && K >= BK, "gemm_make_hilo: K=%d must be a multiple of 64", K);
  int n_tile = 0;
  if (n_tile_req > 0) { if (n_tile_req % 16 == 0 && n_tile_req <= 256 && N % n_tile_req == 0) n_tile = n_tile_req; }
  else for (int c = 256; c >= 16; c -= 16) if (N % c == 0) { n_tile = c; break; }
  CBX_CHECK(n_tile > 0, "gemm_make_hilo: N=%d unsupported (tile %d)", N, n_tile_req);
  CBX_CHECK(N_w >= 1 && N_w <= N, "gemm_make_hilo: %d weight rows for %d output columns", N_w, N);
  CBX_CHECK(!epi.swiglu || n_tile == 256, "gemm_make_hilo: SwiGLU pairs the halves of a 256-column tile (N=%d)", N);
  CBX_CHECK(!epi.ln_g && !epi.out_ln, "gemm_make_hilo: no row LayerNorm on this path");
  d->B = 1; d->L = M; d->N = N; d->K = K; d->taps = 1; d->n_tile = n_tile; d->epi = epi;
  d->dil = 0; d->pad = 0; d->M_out = M; d->hilo = 1; d->nsplit = 1;
  const int tiles = ((M + BM - 1) / BM) * (N / n_tile), nkb_all = K / BK;
  d->ksplit = 1;
  if (ksplit_req > 0) {
    CBX_CHECK(ksplit_req <= MAX_SPLIT && (ksplit_req & (ksplit_req - 1)) == 0 && nkb_all >= ksplit_req,
              "gemm_make_hilo: split-K %d unsupported", ksplit_req);
    d->ksplit = ksplit_req;
  } else {
    for (int sp = 2; sp >= 2; sp >>= 1)
      if (nkb_all >= sp && tiles * sp <= 148) { d->ksplit = sp; break; }
  }
  if (make_tmap_bf16_3d(&d->tmA, A, K, M, 2, lda, (uint64_t)plane, BK, BM)) return CBX_ERR;
  // N_w < N: weight rows past the matrix arrive as zeros (TMA out-of-bounds fill), the output buffer is padded to N
  if (make_tmap_bf16_3d(&d->tmW, W, K, N_w, 1, K, (uint64_t)K * N_w, BK, n_tile)) return CBX_ERR;
  return CBX_OK;
}

bool pdl_enabled() {
  static const bool on = [] { const char* v = getenv("CBX_PDL"); return !(v && v[0] == '0'); }();
  return on;
}

static unsigned long long* g_gemm_dbg = nullptr;
void gemm_set_debug(unsigned long long* buf) { g_gemm_dbg = buf; }

namespace {
uint32_t epi_features(const GemmEpi& e) {
  uint32_t f = 0;
  if (e.ln_g) f |= F_LN1;
  if (e.gelu) f |= F_GELU;
  if (e.elu) f |= F_ELU;
  if (e.resid) f |= F_RESID;
  if (e.resid2) f |= F_RESID2;
  if (e.out_f32) f |= F_F32;
  if (e.out_bf16) f |= F_BF16;
  if (e.out_ln) f |= F_LN2;
  if (e.snake_alpha || e.lrelu != 0.f) f |= F_ACT;
  if (e.out_limit != 0 || e.out_shift != 0) f |= F_SHIFT;
  if (e.out_hilo) f |= F_HILO;
  if (e.swiglu) f |= F_SWIGLU;
  return f;
}
template <uint32_t FEAT>
int launch_as(const GemmParams& p, dim3 grid, cudaStream_t st) {
  static CbxPerDeviceOnce once;        // per device: a single-process multi-GPU server launches on every ordinal
  if (once.first())
    CBX_CUDA(cudaFuncSetAttribute(s3_gemm_kernel<FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = dim3(GEMM_THREADS); cfg.dynamicSmemBytes = GEMM_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (grid.z > 1 || p.nsplit > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 1; at[na].val.clusterDim.y = p.nsplit; at[na].val.clusterDim.z = grid.z;
    ++na;
  }
  if (pdl_enabled()) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at; cfg.numAttrs = na;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_gemm_kernel<FEAT>, p));
  return CBX_OK;
}
// instantiations, narrowest first: the estimator's five epilogue shapes, then everything
constexpr uint32_t V_PLAIN = F_BF16;                                  // q/k/v projection
constexpr uint32_t V_GELU = F_BF16 | F_GELU;                          // feed-forward up-projection
constexpr uint32_t V_RESLN = F_RESID | F_F32 | F_LN2 | F_BF16;        // attention out / feed-forward down (+ next LayerNorm)
constexpr uint32_t V_LN1 = F_LN1 | F_BF16;                            // resnet block1: conv + LayerNorm + Mish + time embedding
constexpr uint32_t V_LN1R = F_LN1 | F_RESID | F_F32 | F_LN2 | F_BF16; // resnet block2 (+ skip path, + next LayerNorm)
constexpr uint32_t V_T3_F32 = F_RESID | F_F32;                        // T3 prefill: q|k|v (plain fp32) / o_proj, down (+ residual)
constexpr uint32_t V_T3_GU = F_SWIGLU | F_HILO;                       // T3 prefill: gate/up + SwiGLU -> hi/lo planes
}  // namespace

bool gemm_ws_eligible(const GemmDesc& d);
int gemm_ws_launch(const GemmDesc& d, cudaStream_t st);

int gemm_launch(const GemmDesc& d, cudaStream_t st) {
  if (gemm_ws_eligible(d)) return gemm_ws_launch(d, st);      // padded batches: weight-stationary persistent kernel (s3_gemm_ws.cu)
  GemmParams p;
  p.tmA = d.tmA; p.tmW = d.tmW;
  p.B = d.B; p.L = d.L; p.N = d.N; p.K = d.K; p.taps = d.taps; p.n_tile = d.n_tile;
  p.dil = d.dil; p.pad = d.pad; p.M_out = d.M_out; p.ksplit = d.ksplit; p.hilo = d.hilo; p.nsplit = d.nsplit;
  p.tiles_l = (d.M_out + BM - 1) / BM;
  p.stage_bytes = ((d.hilo ? 2 : 1) * A_BYTES + d.n_tile * BK * 2 + 1023) / 1024 * 1024;   // swizzled tiles start 1024-byte aligned
  p.nstages = (STAGES * STAGE_BYTES) / p.stage_bytes;
  if (p.nstages > MAX_STAGES) p.nstages = MAX_STAGES;
  p.e = d.epi;
  p.dbg = g_gemm_dbg;
  if (g_gemm_dbg) g_gemm_dbg += 16;
  const dim3 grid(d.B * p.tiles_l, d.N / d.n_tile, d.ksplit);
  const uint32_t f = epi_features(d.epi);
  auto within = [f](uint32_t v) { return (f & ~v) == 0; };
  if (within(V_PLAIN)) return launch_as<V_PLAIN>(p, grid, st);
  if (within(V_T3_F32)) return launch_as<V_T3_F32>(p, grid, st);
  if (within(V_T3_GU)) return launch_as<V_T3_GU>(p, grid, st);
  if (within(V_GELU)) return launch_as<V_GELU>(p, grid, st);
  if (within(V_RESLN)) return launch_as<V_RESLN>(p, grid, st);
  if (within(V_LN1)) return launch_as<V_LN1>(p, grid, st);
  if (within(V_LN1R)) return launch_as<V_LN1R>(p, grid, st);
  return launch_as<F_ALL>(p, grid, st);
}

}  // namespace cbx
