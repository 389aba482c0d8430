// tcgen05 flash attention for the S3Gen estimator's BasicTransformerBlock (reference: matcha/transformer.py:196-204 ->
// diffusers Attention / AttnProcessor2_0: 8 heads x 64, scale 1/8, bidirectional, key-padding mask only
// (decoder.py:25-33, utils/mask.py:167-193)).
//
// One CTA owns 128 query rows of one (batch row, head).  Keys are taken in blocks of 128:
//   warp 0      TMA producer: Q tile once, then K block [128 keys][64 d] and V block [128 keys][64 d] per stage
//               (3-stage ring, 128-byte swizzle), straight out of the q|k|v buffer the GEMM wrote;
//   warp 1      MMA issuer:  S = Q K^T   (tcgen05.mma M128 N128 K16 x 4, both operands K-major)  -> TMEM S[g]
//                            O~ = P V     (M128 N64 K16 x 8; A = P from shared memory, B = the V block as it lies in
//                                          memory = an MN-major operand: no transpose anywhere) -> TMEM O[g]
//   warps 2-5 / 6-9   two softmax groups taking alternate key blocks (g = block & 1), thread <-> query row (TMEM lane):
//               two passes over the S tile in TMEM (row max, then exp2 / row sum / bf16 P written to shared memory in the
//               K-major swizzled operand layout), the running output lives in registers: o = o * corr + O~ with O~ read
//               back from TMEM one block late, so the P V product of a block overlaps the softmax of the next.
// The two groups' (m, l, o) states are merged through shared memory at the end (as the mma.sync kernel did).
// fp32 scores, fp32 online softmax, bf16 P (as before).  Replaces the mma.sync kernel of csrc/s3_attn.cu (kept: CBX_ATTN_TC=0).
#include "s3_gemm.cuh"
#include <cstdlib>

namespace cbx {
namespace {

constexpr int TQ = 128, TK = 128, TD = 64;
constexpr int NST = 3;                                 // K/V stages
constexpr int TILE_B = TQ * TD * 2;                    // 16 KB: [128][64] bf16
constexpr int STAGE_B = 2 * TILE_B;                    // K + V
constexpr int P_B = TQ * TK * 2;                       // 32 KB: two k-blocks of [128][64]
constexpr int OFF_Q = 0, OFF_ST = TILE_B, OFF_P = OFF_ST + NST * STAGE_B, OFF_ML = OFF_P + 2 * P_B;
constexpr int OFF_BAR = OFF_ML + TQ * 2 * 4;
constexpr int ATC_THREADS = 320;
constexpr int ATC_SMEM = OFF_BAR + 256 + 1024;
static_assert(TQ * TD * 4 <= P_B, "group-1 output state fits its P buffer");

struct AttnTcParams {
  CUtensorMap tm;            // q|k|v bf16 [B][L][1536], box {64, 128, 1}
  __nv_bfloat16* out;        // [B][L][512]
  const int* lens;
  int L;
};

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
  unsigned spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void group_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

__global__ void __launch_bounds__(ATC_THREADS, 1) s3_attn_tc_kernel(const __grid_constant__ AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem + OFF_Q;
  uint8_t* sP = smem + OFF_P;
  float* sML = reinterpret_cast<float*>(smem + OFF_ML);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_q = bars;
  uint64_t* full = bars + 1;             // [NST]
  uint64_t* empty = full + NST;          // [NST]
  uint64_t* s_full = empty + NST;        // [2]
  uint64_t* s_free = s_full + 2;         // [2]
  uint64_t* p_full = s_free + 2;         // [2]
  uint64_t* o_full = p_full + 2;         // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * TQ, head = blockIdx.y, b = blockIdx.z;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    mbar_init(bar_q, 1);
    for (int i = 0; i < NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 128); mbar_init(&p_full[i], 128); mbar_init(&o_full[i], 1); }
    mbar_fence_init();
    tma_prefetch_desc(&p.tm);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int Lk = p.lens ? min(max(p.lens[b], 1), p.L) : p.L;
  const int nkb = (Lk + TK - 1) / TK;

  if (warp == 0) {
    if (lane == 0) {
      pdl_wait();                                  // q|k|v come from the GEMM just before
      mbar_arrive_expect_tx(bar_q, TILE_B);
      tma_load_3d(sQ, &p.tm, bar_q, head * TD, q0, b);
      for (int j = 0; j < nkb; ++j) {
        const int s = j % NST;
        if (j >= NST) mbar_wait_sleep(&empty[s], ((j / NST) - 1) & 1);
        uint8_t* st = smem + OFF_ST + s * STAGE_B;
        mbar_arrive_expect_tx(&full[s], STAGE_B);
        tma_load_3d(st, &p.tm, &full[s], 512 + head * TD, j * TK, b);
        tma_load_3d(st + TILE_B, &p.tm, &full[s], 1024 + head * TD, j * TK, b);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16(TQ, TK);
      const uint32_t idesc_o = umma_idesc_bf16(TQ, TD) | (1u << 16);        // B (= V) is MN-major
      mbar_wait_sleep(bar_q, 0);
      for (int j = 0; j <= nkb; ++j) {
        if (j < nkb) {
          const int g = j & 1, it = j >> 1, s = j % NST;
          mbar_wait_sleep(&full[s], (j / NST) & 1);
          if (it > 0) mbar_wait_sleep(&s_free[g], (it - 1) & 1);
          tc_fence_after();
          const uint64_t da = umma_desc_k128(smem_u32(sQ)), db = umma_desc_k128(smem_u32(smem + OFF_ST + s * STAGE_B));
#pragma unroll
          for (int k = 0; k < TD / 16; ++k) umma_bf16(tmem_base + g * TK, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_s, k != 0);
          umma_commit(&s_full[g]);
        }
        if (j >= 1) {
          const int jj = j - 1, g = jj & 1, it = jj >> 1, s = jj % NST;
          mbar_wait_sleep(&p_full[g], it & 1);
          tc_fence_after();
          const uint32_t pa = smem_u32(sP + g * P_B), va = smem_u32(smem + OFF_ST + s * STAGE_B + TILE_B);
#pragma unroll
          for (int k = 0; k < TK / 16; ++k) {
            // A: k-block k / 4 of P, 32 bytes per K = 16 step inside the swizzle row; B: 16 key rows = 2048 bytes per step
            const uint64_t da = umma_desc_k128(pa + (k >> 2) * TILE_B) + (uint64_t)((k & 3) * 2);
            const uint64_t db = umma_desc_k128(va + k * 2048);
            umma_bf16(tmem_base + 2 * TK + g * TD, da, db, idesc_o, k != 0);
          }
          umma_commit(&o_full[g]);
          umma_commit(&empty[s]);
        }
      }
    }
    __syncwarp();
  } else {
    // ---- softmax groups ----
    const int g = (warp - 2) >> 2, q = warp & 3, row = q * 32 + lane;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t tS = tmem_base + lane_off + g * TK, tO = tmem_base + lane_off + 2 * TK + g * TD;
    const float sc = 0.125f * 1.4426950408889634f;              // 1/sqrt(64) * log2(e): scores in the log2 domain
    float m = -INFINITY, l = 0.f, corr_prev = 1.f;
    float o[TD];
#pragma unroll
    for (int d = 0; d < TD; ++d) o[d] = 0.f;
    uint8_t* prow = sP + g * P_B + row * 128;
    int i = 0;
    for (int j = g; j < nkb; j += 2, ++i) {
      mbar_wait_sleep(&s_full[g], i & 1);
      tc_fence_after();
      const int kbase = j * TK;
      const bool tail = kbase + TK > Lk;
      // pass 1: row maximum
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld32(tS + c * 32, v);
        if (tail) {
#pragma unroll
          for (int k = 0; k < 32; ++k) if (kbase + c * 32 + k >= Lk) v[k] = -INFINITY;
        }
#pragma unroll
        for (int k = 0; k < 32; ++k) mx = fmaxf(mx, v[k]);
      }
      const float m_new = fmaxf(m, mx * sc);
      const float corr = ex2f(m - m_new);           // first block: exp2(-inf) = 0, and o = l = 0
      // fold the previous block's P V product into the running output (it has been computing meanwhile); this also
      // says the P buffer is free again
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
      corr_prev = corr;
      // pass 2: p = exp2(s - m), row sum, bf16 P in the swizzled K-major layout
      float ls = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float v[32];
        tmem_ld32(tS + c * 32, v);
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          float pv = ex2f(fmaf(v[k], sc, -m_new));
          if (tail && kbase + c * 32 + k >= Lk) pv = 0.f;
          v[k] = pv; ls += pv;
        }
        uint8_t* pk = prow + (c >> 1) * TILE_B;
#pragma unroll
        for (int jc = 0; jc < 4; ++jc) {
          const uint4 w = make_uint4(pack_bf16x2(v[jc * 8], v[jc * 8 + 1]), pack_bf16x2(v[jc * 8 + 2], v[jc * 8 + 3]),
                                     pack_bf16x2(v[jc * 8 + 4], v[jc * 8 + 5]), pack_bf16x2(v[jc * 8 + 6], v[jc * 8 + 7]));
          *reinterpret_cast<uint4*>(pk + ((((c & 1) * 4 + jc) ^ (row & 7)) << 4)) = w;
        }
      }
      l = l * corr + ls;
      m = m_new;
      tc_fence_before();
      mbar_arrive(&s_free[g]);                                          // S[g] may be overwritten
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // P visible to the tensor core
      mbar_arrive(&p_full[g]);
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
    tc_fence_before();
    // ---- merge group 1 into group 0 (group 1's P buffer is idle: its last product has completed) ----
    float* xo = reinterpret_cast<float*>(sP + P_B);           // [64 d][128 rows]
    if (g == 1) {
#pragma unroll
      for (int d = 0; d < TD; ++d) xo[d * TQ + row] = o[d];
      sML[row * 2] = m; sML[row * 2 + 1] = l;
    }
    asm volatile("bar.sync 2, 256;" ::: "memory");
    if (g == 0) {
      const float m1 = sML[row * 2], l1 = sML[row * 2 + 1];
      const float mn = fmaxf(m, m1);
      const float c0 = (m == -INFINITY) ? 0.f : ex2f(m - mn), c1 = (m1 == -INFINITY) ? 0.f : ex2f(m1 - mn);
      const float inv = 1.0f / (l * c0 + l1 * c1);
      const int t = q0 + row;
      if (t < p.L) {
        __nv_bfloat16* dst = p.out + ((size_t)b * p.L + t) * 512 + head * TD;
#pragma unroll
        for (int d8 = 0; d8 < TD / 8; ++d8) {
          float y[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) y[e] = (o[d8 * 8 + e] * c0 + xo[(d8 * 8 + e) * TQ + row] * c1) * inv;
          *reinterpret_cast<uint4*>(dst + d8 * 8) = make_uint4(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]),
                                                               pack_bf16x2(y[4], y[5]), pack_bf16x2(y[6], y[7]));
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

int s3_attn_tc_launch(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st) {
  CBX_CHECK(ld_qkv == 1536, "s3_attn_tc: q|k|v row stride %d (expected 1536)", ld_qkv);
  static CbxPerDeviceOnce once;
  if (once.first()) CBX_CUDA(cudaFuncSetAttribute(s3_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATC_SMEM));
  AttnTcParams p;
  if (make_tmap_bf16_3d(&p.tm, qkv, 1536, L, B, 1536, (uint64_t)1536 * L, TD, TQ)) return CBX_ERR;
  p.out = out; p.lens = lens; p.L = L;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((L + TQ - 1) / TQ, 8, B); cfg.blockDim = dim3(ATC_THREADS); cfg.dynamicSmemBytes = ATC_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_attn_tc_kernel, p));
  return CBX_OK;
}

}  // namespace cbx
