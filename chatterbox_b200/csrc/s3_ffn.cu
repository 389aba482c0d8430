// Fused feed-forward of the S3Gen estimator's BasicTransformerBlock (reference: matcha/transformer.py:300-316 —
// norm3 -> FeedForward(GELU, 256 -> 1024 -> 256) -> residual; the LayerNorm that follows (the next block's norm1,
// transformer.py:262) is folded in as in the GEMM epilogue).
//
// At the estimator's sizes (2 x 812 rows = 14 row tiles of 128) the two Linear layers as separate tcgen05 GEMM launches
// are latency-bound: 13 + 16 us for 1.7 GFLOP.  Here ONE cluster of 8 CTAs owns a 128-row tile; CTA c of the cluster owns
// hidden columns [128 c, 128 c + 128):
//   TMA      x tile [128][256] + W1 rows [128 c, +128) x [256]  (4 k-blocks, one mbarrier each), W2[:, 128 c .. +128)
//            — the weight panels are requested before the programmatic-launch wait (they never depend on a kernel);
//   MMA 1    D1[128][128] = x W1c^T           (tcgen05.mma M128 N128 K16 x 16, TMEM columns 0..127)
//   epi 1    D1 + b1 -> GELU -> bf16 -> shared memory in the canonical K-major 128-byte-swizzle layout (the A operand
//            of the second product), fence.proxy.async
//   MMA 2    D2[128][256] = h W2c^T           (M128 N256 K16 x 8, TMEM columns 128..383)
//   reduce   D2 -> fp32 partial tile in shared memory, cluster barrier, CTA c sums rows [16 c, 16 c + 16) of the 8
//            partial tiles through distributed shared memory (rank order: deterministic), + b2 + residual -> fp32,
//            LayerNorm -> bf16 for the next GEMM.
// The hidden activations never leave the SM; 112 CTAs work on what 14 row tiles offer.
#include "s3_gemm.cuh"
#include <cstdlib>

namespace cbx {
namespace {

constexpr int FC = 8;                    // CTAs per cluster = hidden chunks
constexpr int FH = 1024, FD = 256;       // hidden / model width
constexpr int HC = FH / FC;              // 128 hidden columns per CTA
constexpr int FBM = 128, FBK = 64;
constexpr int FFN_THREADS = 512;
constexpr int X_BYTES = FBM * FD * 2;            // 64 KB: 4 k-blocks of [128][64]
constexpr int W1_BYTES = HC * FD * 2;            // 64 KB: 4 k-blocks of [128][64]
constexpr int W2_BYTES = FD * HC * 2;            // 64 KB: 2 k-blocks of [256][64]
constexpr int KB_A = FBM * FBK * 2;              // 16 KB
constexpr int KB_W2 = FD * FBK * 2;              // 32 KB
constexpr int PITCH = FD + 4;                    // floats per row of the partial tile
constexpr int PAR_OFF = X_BYTES + W1_BYTES + W2_BYTES;        // b1[128], b2[256], g[256], b[256]
constexpr int OC = FD / FC;                      // 32 out-projection columns per CTA (fused attention out-projection)
constexpr int AK = 512, A_KB = AK / FBK;         // attention width = K of the out-projection, 8 k-blocks
constexpr int WO_KB = OC * FBK * 2;              // 4 KB: [32 rows][64 k]
constexpr int PAR_N = HC + 3 * FD + 3 * OC;      // + bo[32], g0[32], b0[32] of the fused out-projection
constexpr int STAT_OFF = PAR_OFF + 4096;         // [8 senders][128 rows] (sum, sum of squares)
constexpr int BAR_OFF = STAT_OFF + FC * FBM * 8;
constexpr int FFN_SMEM = BAR_OFF + 256 + 1024 /*align*/;
static_assert(A_KB == FC, "one attention k-block per CTA of the cluster (multicast)");
static_assert(PAR_N * 4 <= 4096 && A_KB * KB_A <= X_BYTES + W1_BYTES && A_KB * WO_KB <= X_BYTES, "fused out-projection buffers");
static_assert(FBM * PITCH * 4 <= PAR_OFF, "partial tile overlays the operand buffers");
static_assert(FFN_SMEM <= 232448, "shared memory budget");

struct FfnParams {
  CUtensorMap tmX, tmW1, tmW2;
  int L, tiles_l;
  const float *b1, *b2, *resid, *ln_g, *ln_b;
  float* out_f32;
  __nv_bfloat16 *out_ln, *out_bf16;
  int ld_resid, ld_f32, ld_ln, ld_bf16;
  // fused attention out-projection (OUT instantiation): x1 = resid0 + att Wo^T + bo -> out0_f32; LayerNorm(x1) -> xbuf (= the
  // tensor tmX reads)
  CUtensorMap tmA, tmWo;
  const float *bo, *g0, *b0, *resid0;
  float* out0_f32;
  __nv_bfloat16* xbuf;
  int ld_res0, ld_out0, ld_x;
  float* red;                  // optional global scratch [row tiles][8][128][256] fp32: the partial tiles meet in L2 instead of DSMEM
  unsigned long long* dbg;    // optional %globaltimer stamps of CTA (0, 0, 0) (debug aid, tools/ffn_stamps.py)
};

__device__ __forceinline__ unsigned long long ffn_gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define FFN_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && blockIdx.z == 0) p.dbg[k] = ffn_gtimer(); } while (0)

template <bool OUT>
__global__ void __cluster_dims__(1, 1, FC) __launch_bounds__(FFN_THREADS, 1) s3_ffn_kernel(const __grid_constant__ FfnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  // plain:  [x 64 KB | W1 slice 64 KB | W2 slice 64 KB]
  // OUT:    [attention tile 128 KB -> W1 slice | W2 slice] [Wo slice 32 KB -> x 64 KB]: the weights of the feed-forward
  //         take the attention tile's place once the out-projection's MMAs have completed
  uint8_t* sX = smem + (OUT ? X_BYTES + W1_BYTES : 0);         // x tile; after MMA 1 its first half holds h
  uint8_t* sW1 = smem + (OUT ? 0 : X_BYTES);
  uint8_t* sW2 = smem + (OUT ? W1_BYTES : X_BYTES + W1_BYTES);
  uint8_t* sA = smem;                                          // OUT: attention tile, 8 k-blocks of [128][64]
  uint8_t* sWo = smem + X_BYTES + W1_BYTES;                    // OUT: Wo rows [32 rank, +32), 8 k-blocks of [32][64]
  float2* stat = reinterpret_cast<float2*>(smem + STAT_OFF);
  float* par = reinterpret_cast<float*>(smem + PAR_OFF);
  uint64_t* bar_k = reinterpret_cast<uint64_t*>(smem + BAR_OFF);    // [4] x + W1 k-block landed
  uint64_t* bar_w2 = bar_k + 4;
  uint64_t* bar_d1 = bar_k + 5;
  uint64_t* bar_d2 = bar_k + 6;
  uint64_t* bar_a = bar_k + 7;                    // OUT: [8] attention + Wo k-block landed
  uint64_t* bar_d0 = bar_k + 15;                  // OUT: out-projection accumulator ready
  uint64_t* stat_bar = bar_k + 16;                // OUT: the 8 CTAs' row moments have landed (st.async bytes)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_k + 17);
  float* tile = reinterpret_cast<float*>(smem);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = blockIdx.z;
  const int b = blockIdx.x / p.tiles_l, t0 = (blockIdx.x % p.tiles_l) * FBM;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    FFN_STAMP(0);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_k[i], 1);
    mbar_init(bar_w2, 1); mbar_init(bar_d1, 1); mbar_init(bar_d2, 1);
    if (OUT) {
      for (int i = 0; i < A_KB; ++i) mbar_init(&bar_a[i], 1);
      mbar_init(bar_d0, 1); mbar_init(stat_bar, 1);
      mbar_arrive_expect_tx(stat_bar, (uint32_t)(FC * FBM * 8));
    }
    mbar_fence_init();
    tma_prefetch_desc(&p.tmX); tma_prefetch_desc(&p.tmW1); tma_prefetch_desc(&p.tmW2);
    if (OUT) { tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmWo); }
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_d1 = tmem_base, tmem_d2 = tmem_base + 128, tmem_d0 = tmem_base + 384;
  // OUT: peers push row moments into our shared memory later: tell them our barriers exist (waited for before the first push)
  if (OUT) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  if (threadIdx.x == 0) FFN_STAMP(1);

  // final-epilogue ownership: this CTA finishes rows [16 rank, 16 rank + 16) of the tile, one row per warp; a lane
  // holds the float4 column groups lane and lane + 32
  const int frow = rank * (FBM / FC) + warp, ft = t0 + frow;
  const bool frow_ok = ft < p.L;
  const size_t grow = (size_t)b * p.L + (frow_ok ? ft : 0);
  const int gc0 = lane * 4, gc1 = (lane + 32) * 4;
  float4 rv0 = make_float4(0.f, 0.f, 0.f, 0.f), rv1 = rv0;

  auto issue_weights = [&]() {          // the feed-forward's weight slices (never depend on a kernel)
    for (int kb = 0; kb < 4; ++kb) {
      mbar_arrive_expect_tx(&bar_k[kb], 2 * KB_A);
      tma_load_3d(sW1 + kb * KB_A, &p.tmW1, &bar_k[kb], kb * FBK, rank * HC, 0);
    }
    mbar_arrive_expect_tx(bar_w2, W2_BYTES);
    for (int kb = 0; kb < 2; ++kb) tma_load_3d(sW2 + kb * KB_W2, &p.tmW2, bar_w2, rank * HC + kb * FBK, 0, 0);
  };
  auto mma1 = [&]() {
    const uint32_t idesc = umma_idesc_bf16(FBM, HC);
    for (int kb = 0; kb < 4; ++kb) {
      mbar_wait_spin(&bar_k[kb], 0);
      if (kb == 0) FFN_STAMP(3);
      if (kb == 3) FFN_STAMP(4);
      tc_fence_after();
      const uint64_t da = umma_desc_k128(smem_u32(sX + kb * KB_A)), db = umma_desc_k128(smem_u32(sW1 + kb * KB_A));
#pragma unroll
      for (int k = 0; k < FBK / 16; ++k) umma_bf16(tmem_d1, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
    }
    umma_commit(bar_d1);
  };
  if (warp == 0) {
    // OUT: the attention tile is the same for the 8 CTAs of the cluster: each CTA fetches ONE of its 8 k-blocks and
    // multicasts it into all 8 shared memories (1 / 8 of the L2 reads).  A peer's barriers must exist before data may land
    // in its shared memory: the producer warp takes the cluster barrier here, the other warps further down.
    if (OUT) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (elect_one()) {
      if (!OUT) {
        issue_weights();
        pdl_wait();
        FFN_STAMP(2);
        for (int kb = 0; kb < 4; ++kb) tma_load_3d(sX + kb * KB_A, &p.tmX, &bar_k[kb], kb * FBK, t0, b);
      } else {
        for (int kb = 0; kb < A_KB; ++kb) {
          mbar_arrive_expect_tx(&bar_a[kb], KB_A + WO_KB);
          tma_load_3d(sWo + kb * WO_KB, &p.tmWo, &bar_a[kb], kb * FBK, rank * OC, 0);
        }
        pdl_wait();
        FFN_STAMP(2);
#ifndef FFN_NO_MULTICAST
        tma_load_3d_multicast(sA + rank * KB_A, &p.tmA, &bar_a[rank], rank * FBK, t0, b, (uint16_t)0xFF);
#else
        for (int kb = 0; kb < A_KB; ++kb) tma_load_3d(sA + kb * KB_A, &p.tmA, &bar_a[kb], kb * FBK, t0, b);
#endif
        mbar_wait_spin(bar_d0, 0);        // the out-projection's MMAs are done with the attention tile: its place is free
        issue_weights();
      }
    }
    __syncwarp();
    pdl_wait();
  } else if (warp == 1) {
    if (elect_one()) {
      if (!OUT) mma1();
      else {
        const uint32_t idesc0 = umma_idesc_bf16(FBM, OC);
        for (int kb = 0; kb < A_KB; ++kb) {
          mbar_wait_spin(&bar_a[kb], 0);
          tc_fence_after();
          const uint64_t da = umma_desc_k128(smem_u32(sA + kb * KB_A)), db = umma_desc_k128(smem_u32(sWo + kb * WO_KB));
#pragma unroll
          for (int k = 0; k < FBK / 16; ++k) umma_bf16(tmem_d0, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc0, (kb | k) != 0);
        }
        umma_commit(bar_d0);
      }
    }
    __syncwarp();
    pdl_wait();
  } else {
    // per-column vectors of the epilogues
    for (int c = threadIdx.x - 64; c < (OUT ? PAR_N : HC + 3 * FD); c += FFN_THREADS - 64) {
      float v;
      if (c < HC) v = __ldg(p.b1 + rank * HC + c);
      else if (c < HC + FD) v = __ldg(p.b2 + c - HC);
      else if (c < HC + 2 * FD) v = p.ln_g ? __ldg(p.ln_g + c - HC - FD) : 1.f;
      else if (c < HC + 3 * FD) v = p.ln_g ? __ldg(p.ln_b + c - HC - 2 * FD) : 0.f;
      else if (c < HC + 3 * FD + OC) v = __ldg(p.bo + rank * OC + c - (HC + 3 * FD));
      else if (c < HC + 3 * FD + 2 * OC) v = __ldg(p.g0 + rank * OC + c - (HC + 3 * FD + OC));
      else v = __ldg(p.b0 + rank * OC + c - (HC + 3 * FD + 2 * OC));
      par[c] = v;
    }
    pdl_wait();
  }
  if (OUT) {
    // ---- fused attention out-projection: this CTA holds columns [32 rank, +32) of x1 = resid + att Wo^T + bo for the
    // tile's 128 rows (warps 4-7, thread <-> row); the LayerNorm over the 256 channels needs the other CTAs' moments:
    // pushed with st.async, then the normalised bf16 slice goes to global memory (the x buffer), and after a cluster
    // barrier every CTA loads the full x tile by TMA — from here on the kernel is the plain feed-forward. ----
    if (warp != 0) asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");        // every peer's barriers exist
    __syncthreads();                                                          // par[] staged
    if (warp >= 4 && warp < 8) {
      const int q = warp & 3, row = q * 32 + lane, t = t0 + row;
      const bool ok = t < p.L;
      const size_t g0row = (size_t)b * p.L + (ok ? t : 0);
      float4 rs[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) rs[i] = ok ? *reinterpret_cast<const float4*>(p.resid0 + g0row * p.ld_res0 + rank * OC + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
      mbar_wait_spin(bar_d0, 0);
      tc_fence_after();
      if (threadIdx.x == 128) FFN_STAMP(12);
      float v[32];
      tmem_ld32(tmem_d0 + ((uint32_t)(q * 32) << 16), v);
      const float* pb = par + HC + 3 * FD;
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[4 * i] += pb[4 * i] + rs[i].x; v[4 * i + 1] += pb[4 * i + 1] + rs[i].y;
        v[4 * i + 2] += pb[4 * i + 2] + rs[i].z; v[4 * i + 3] += pb[4 * i + 3] + rs[i].w;
      }
#pragma unroll
      for (int k = 0; k < 32; ++k) { s1 += v[k]; s2 += v[k] * v[k]; }
      {
        const uint32_t la = smem_u32(&stat[rank * FBM + row]), lb = smem_u32(stat_bar);
#pragma unroll
        for (int j = 0; j < FC; ++j)
          asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                       ::"r"(dsmem_addr(la, (uint32_t)j)), "f"(s1), "f"(s2), "r"(dsmem_addr(lb, (uint32_t)j)) : "memory");
      }
      if (ok) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(p.out0_f32 + g0row * p.ld_out0 + rank * OC + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
      {
        unsigned spins = 0;
        uint32_t done = 0;
        while (!done) {
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(done) : "r"(smem_u32(stat_bar)), "r"(0u) : "memory");
          if (++spins > (1u << 26)) __trap();
        }
      }
      if (threadIdx.x == 128) FFN_STAMP(13);
      s1 = 0.f; s2 = 0.f;
#pragma unroll
      for (int j = 0; j < FC; ++j) { const float2 m2 = stat[j * FBM + row]; s1 += m2.x; s2 += m2.y; }
      const float mean = s1 * (1.0f / FD);
      const float rstd = rsqrtf(fmaxf(s2 * (1.0f / FD) - mean * mean, 0.f) + 1e-5f);
      if (ok) {
        const float* pg = pb + OC; const float* pbb = pb + 2 * OC;
        uint32_t w[16];
#pragma unroll
        for (int k = 0; k < 16; ++k)
          w[k] = pack_bf16x2((v[2 * k] - mean) * rstd * pg[2 * k] + pbb[2 * k], (v[2 * k + 1] - mean) * rstd * pg[2 * k + 1] + pbb[2 * k + 1]);
        __nv_bfloat16* dst = p.xbuf + g0row * p.ld_x + rank * OC;
#pragma unroll
        for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(dst + 8 * i) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
      }
      asm volatile("fence.proxy.async;" ::: "memory");      // these global writes are read by the peers' TMA loads
    }
    if (threadIdx.x == 128) FFN_STAMP(14);
    tc_fence_before();
    cluster_sync_all();                 // x1 and the normalised x tile of all 8 column slices are in global memory
    tc_fence_after();
    if (threadIdx.x == 128) FFN_STAMP(15);
    if (warp == 0 && elect_one()) {
      asm volatile("fence.proxy.async;" ::: "memory");
      for (int kb = 0; kb < 4; ++kb) tma_load_3d(sX + kb * KB_A, &p.tmX, &bar_k[kb], kb * FBK, t0, b);
    }
    if (warp == 1 && elect_one()) mma1();
    __syncwarp();
  }
  if (frow_ok) {
    rv0 = __ldcg(reinterpret_cast<const float4*>(p.resid + grow * p.ld_resid + gc0));
    rv1 = __ldcg(reinterpret_cast<const float4*>(p.resid + grow * p.ld_resid + gc1));
  }
  // ---- epilogue 1 (all 16 warps): D1 -> + b1 -> GELU -> bf16 -> h in the K-major swizzled operand layout ----
  if (warp >= 2 && warp <= 5) mbar_wait_spin(bar_d1, 0);
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 64) FFN_STAMP(5);
  {
    const int q = warp & 3, row = q * 32 + lane, cc = warp >> 2;          // 32 columns [32 cc, 32 cc + 32)
    float v[32];
    tmem_ld32(tmem_d1 + ((uint32_t)(q * 32) << 16) + cc * 32, v);
    const float* b1 = par + cc * 32;
    uint8_t* hrow = sX + (cc >> 1) * KB_A + row * 128;
    const int j0 = (cc & 1) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t w[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int c = j * 8 + e * 2;
        w[e] = pack_bf16x2(gelu_f(v[c] + b1[c]), gelu_f(v[c + 1] + b1[c + 1]));
      }
      *reinterpret_cast<uint4*>(hrow + (((j0 + j) ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy writes of h -> visible to the MMA
  tc_fence_before();
  __syncthreads();
  // ---- MMA 2 ----
  if (warp == 1 && elect_one()) {
    FFN_STAMP(6);
    tc_fence_after();
    mbar_wait_spin(bar_w2, 0);
    const uint32_t idesc = umma_idesc_bf16(FBM, FD);
    for (int kb = 0; kb < 2; ++kb) {
      const uint64_t da = umma_desc_k128(smem_u32(sX + kb * KB_A)), db = umma_desc_k128(smem_u32(sW2 + kb * KB_W2));
#pragma unroll
      for (int k = 0; k < FBK / 16; ++k) umma_bf16(tmem_d2, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
    }
    umma_commit(bar_d2);
  }
  __syncwarp();
  if (warp >= 2 && warp <= 5) mbar_wait_spin(bar_d2, 0);
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 64) FFN_STAMP(7);
  // ---- drain D2 to the fp32 partial tile (all operand buffers are idle now) ----
  {
    const int q = warp & 3, row = q * 32 + lane, cq = warp >> 2;          // 64 columns [64 cq, 64 cq + 64)
    const uint32_t trow = tmem_d2 + ((uint32_t)(q * 32) << 16) + cq * 64;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v[16];
      tmem_ld16(trow + c * 16, v);
      float4* dst = reinterpret_cast<float4*>(tile + (size_t)row * PITCH + cq * 64 + c * 16);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
  tc_fence_before();
  float* red = p.red ? p.red + (size_t)blockIdx.x * FC * FBM * FD : nullptr;       // this row tile's [8][128][256]
  if (red) {
    // partial tile -> global scratch (L2), coalesced: warp <-> 8 rows, lanes across the row.  The SM-to-SM network moves
    // ~20 B/clk per SM, L2 64 B/clk: at 112 KB per CTA the exchange through L2 is the faster one, and nobody reads our shared
    // memory afterwards, so there is no barrier before exit.
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < FBM / 16; ++rr) {
      const int r = warp * (FBM / 16) + rr;
      float* dst = red + ((size_t)rank * FBM + r) * FD;
      *reinterpret_cast<float4*>(dst + gc0) = *reinterpret_cast<const float4*>(tile + (size_t)r * PITCH + gc0);
      *reinterpret_cast<float4*>(dst + gc1) = *reinterpret_cast<const float4*>(tile + (size_t)r * PITCH + gc1);
    }
  }
  if (threadIdx.x == 64) FFN_STAMP(8);
  cluster_sync_all();
  if (threadIdx.x == 64) FFN_STAMP(9);
  // ---- reduce over the cluster + final epilogue: one row per warp ----
  {
    float4 u0[FC], u1[FC];
    if (red) {
#pragma unroll
      for (int j = 0; j < FC; ++j) {
        u0[j] = __ldcg(reinterpret_cast<const float4*>(red + ((size_t)j * FBM + frow) * FD + gc0));
        u1[j] = __ldcg(reinterpret_cast<const float4*>(red + ((size_t)j * FBM + frow) * FD + gc1));
      }
    } else {
      const uint32_t la0 = smem_u32(tile + (size_t)frow * PITCH + gc0), la1 = smem_u32(tile + (size_t)frow * PITCH + gc1);
#pragma unroll
      for (int j = 0; j < FC; ++j) { u0[j] = dsmem_ld4(dsmem_addr(la0, (uint32_t)j)); u1[j] = dsmem_ld4(dsmem_addr(la1, (uint32_t)j)); }
    }
    float x[8];
    {
      float4 a = u0[0], c = u1[0];
#pragma unroll
      for (int j = 1; j < FC; ++j) {
        a.x += u0[j].x; a.y += u0[j].y; a.z += u0[j].z; a.w += u0[j].w;
        c.x += u1[j].x; c.y += u1[j].y; c.z += u1[j].z; c.w += u1[j].w;
      }
      const float4 ba = *reinterpret_cast<const float4*>(par + HC + gc0), bc = *reinterpret_cast<const float4*>(par + HC + gc1);
      x[0] = a.x + ba.x + rv0.x; x[1] = a.y + ba.y + rv0.y; x[2] = a.z + ba.z + rv0.z; x[3] = a.w + ba.w + rv0.w;
      x[4] = c.x + bc.x + rv1.x; x[5] = c.y + bc.y + rv1.y; x[6] = c.z + bc.z + rv1.z; x[7] = c.w + bc.w + rv1.w;
    }
    if (frow_ok) {
      if (p.out_f32) {
        *reinterpret_cast<float4*>(p.out_f32 + grow * p.ld_f32 + gc0) = make_float4(x[0], x[1], x[2], x[3]);
        *reinterpret_cast<float4*>(p.out_f32 + grow * p.ld_f32 + gc1) = make_float4(x[4], x[5], x[6], x[7]);
      }
      if (p.out_bf16) {
        *reinterpret_cast<uint2*>(p.out_bf16 + grow * p.ld_bf16 + gc0) = make_uint2(pack_bf16x2(x[0], x[1]), pack_bf16x2(x[2], x[3]));
        *reinterpret_cast<uint2*>(p.out_bf16 + grow * p.ld_bf16 + gc1) = make_uint2(pack_bf16x2(x[4], x[5]), pack_bf16x2(x[6], x[7]));
      }
    }
    if (p.out_ln) {
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) { s1 += x[k]; s2 += x[k] * x[k]; }
      s1 = warp_sum(s1); s2 = warp_sum(s2);
      const float mean = s1 * (1.0f / FD);
      const float rstd = rsqrtf(fmaxf(s2 * (1.0f / FD) - mean * mean, 0.f) + 1e-5f);
      if (frow_ok) {
        const float* g = par + HC + FD; const float* bb = par + HC + 2 * FD;
        float y[8];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          y[k] = (x[k] - mean) * rstd * g[gc0 + k] + bb[gc0 + k];
          y[4 + k] = (x[4 + k] - mean) * rstd * g[gc1 + k] + bb[gc1 + k];
        }
        *reinterpret_cast<uint2*>(p.out_ln + grow * p.ld_ln + gc0) = make_uint2(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]));
        *reinterpret_cast<uint2*>(p.out_ln + grow * p.ld_ln + gc1) = make_uint2(pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
      }
    }
  }
  if (threadIdx.x == 64) FFN_STAMP(10);
  if (red) __syncthreads(); else cluster_sync_all();          // DSMEM form: peers may still be reading this CTA's partial tile
  if (threadIdx.x == 64) FFN_STAMP(11);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

static unsigned long long* g_ffn_dbg = nullptr;
void ffn_set_debug(unsigned long long* buf) { g_ffn_dbg = buf; }

int ffn_make(FfnDesc* d, const __nv_bfloat16* X, int ldx, const __nv_bfloat16* W1, const __nv_bfloat16* W2, int B, int L) {
  d->B = B; d->L = L; d->fuse_out = 0; d->red = nullptr;
  d->bo = d->g0 = d->b0 = d->resid0 = nullptr; d->out0_f32 = nullptr; d->xbuf = nullptr; d->ld_res0 = d->ld_out0 = d->ld_x = 0;
  if (make_tmap_bf16_3d(&d->tmX, X, FD, L, B, ldx, (uint64_t)ldx * L, FBK, FBM)) return CBX_ERR;
  if (make_tmap_bf16_3d(&d->tmW1, W1, FD, FH, 1, FD, (uint64_t)FD * FH, FBK, HC)) return CBX_ERR;
  if (make_tmap_bf16_3d(&d->tmW2, W2, FH, FD, 1, FH, (uint64_t)FH * FD, FBK, FD)) return CBX_ERR;
  return CBX_OK;
}

int ffn_make_out(FfnDesc* d, const __nv_bfloat16* att, int lda, const __nv_bfloat16* Wo) {
  if (make_tmap_bf16_3d(&d->tmA, att, AK, d->L, d->B, lda, (uint64_t)lda * d->L, FBK, FBM)) return CBX_ERR;
  if (make_tmap_bf16_3d(&d->tmWo, Wo, AK, FD, 1, AK, (uint64_t)AK * FD, FBK, OC)) return CBX_ERR;
  d->fuse_out = 1;
  return CBX_OK;
}

template <bool OUT>
static int ffn_launch_as(const FfnParams& p, dim3 grid, cudaStream_t st) {
  static CbxPerDeviceOnce once;
  if (once.first()) CBX_CUDA(cudaFuncSetAttribute(s3_ffn_kernel<OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, FFN_SMEM));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = dim3(FFN_THREADS); cfg.dynamicSmemBytes = FFN_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_ffn_kernel<OUT>, p));
  return CBX_OK;
}

int ffn_launch(const FfnDesc& d, cudaStream_t st) {
  FfnParams p;
  p.tmX = d.tmX; p.tmW1 = d.tmW1; p.tmW2 = d.tmW2;
  p.L = d.L; p.tiles_l = (d.L + FBM - 1) / FBM;
  p.b1 = d.b1; p.b2 = d.b2; p.resid = d.resid; p.ln_g = d.ln_g; p.ln_b = d.ln_b;
  p.out_f32 = d.out_f32; p.out_ln = d.ln_g ? d.out_ln : nullptr; p.out_bf16 = d.out_bf16;
  p.ld_resid = d.ld_resid; p.ld_f32 = d.ld_f32; p.ld_ln = d.ld_ln; p.ld_bf16 = d.ld_bf16;
  p.tmA = d.tmA; p.tmWo = d.tmWo; p.bo = d.bo; p.g0 = d.g0; p.b0 = d.b0; p.resid0 = d.resid0; p.out0_f32 = d.out0_f32;
  p.xbuf = d.xbuf; p.ld_res0 = d.ld_res0; p.ld_out0 = d.ld_out0; p.ld_x = d.ld_x;
  p.red = d.red;
  p.dbg = g_ffn_dbg;
  if (g_ffn_dbg) g_ffn_dbg += 16;
  const dim3 grid(d.B * p.tiles_l, 1, FC);
  return d.fuse_out ? ffn_launch_as<true>(p, grid, st) : ffn_launch_as<false>(p, grid, st);
}

}  // namespace cbx
