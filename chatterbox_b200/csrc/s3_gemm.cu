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
//   warps 2-5: epilogue — each thread owns one accumulator row (TMEM lane) and walks its columns with
//              tcgen05.ld: bias, LayerNorm-over-channels + Mish (+ time embedding), exact GELU,
//              residual add, fp32 / bf16 stores, and the LayerNorm feeding the next GEMM.
#include "s3_gemm.cuh"
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

constexpr int BM = 128, BK = 64, STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;                 // 16 KB
constexpr int W_BYTES_MAX = 256 * BK * 2;            // 32 KB
constexpr int STAGE_BYTES = A_BYTES + W_BYTES_MAX;
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;

struct GemmParams {
  CUtensorMap tmA, tmW;
  int B, L, N, K, taps, n_tile, tiles_l, dil, pad, M_out;
  GemmEpi e;
  unsigned long long* dbg;    // optional %globaltimer stamps of CTA (0,0) (debug aid)
};

__device__ __forceinline__ float mish_f(float x) {
  const float sp = (x > 20.f) ? x : log1pf(expf(x));
  return x * tanhf(sp);
}
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define GEMM_STAMP(k) do { if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0) p.dbg[k] = gtimer(); } while (0)

__global__ void __launch_bounds__(GEMM_THREADS, 1) s3_gemm_kernel(const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* accum_bar = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) GEMM_STAMP(0);
  const int b = blockIdx.x / p.tiles_l, t0 = (blockIdx.x % p.tiles_l) * BM;
  const int n0 = blockIdx.y * p.n_tile;
  const int kpb = p.K / BK, nkb = p.taps * kpb;
  const uint32_t tmem_cols = p.n_tile > 128 ? 256u : (p.n_tile > 64 ? 128u : (p.n_tile > 32 ? 64u : 32u));

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmW);
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) GEMM_STAMP(1);

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t tx = A_BYTES + p.n_tile * BK * 2;
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES, ph = (kb / STAGES) & 1;
        mbar_wait_spin(&empty[s], ph ^ 1);
        const int tap = kb / kpb, kk = kb % kpb;
        uint8_t* sa = smem + s * STAGE_BYTES;
        mbar_arrive_expect_tx(&full[s], tx);
        tma_load_3d(sa, &p.tmA, &full[s], kk * BK, t0 + tap * p.dil - p.pad, b);
        tma_load_3d(sa + A_BYTES, &p.tmW, &full[s], kb * BK, n0, 0);
      }
      GEMM_STAMP(2);
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(BM, p.n_tile);
      for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES, ph = (kb / STAGES) & 1;
        mbar_wait_spin(&full[s], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
        const uint64_t da = umma_desc_k128(sa), db = umma_desc_k128(sa + A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)    // +32 bytes per K=16 step inside the 128-byte swizzle row
          umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
        umma_commit(&empty[s]);
      }
      umma_commit(accum_bar);
      GEMM_STAMP(3);
    }
  } else {
    // ===== epilogue =====
    // phase 1: thread <-> accumulator row (TMEM lane): bias, LayerNorm + Mish (+ time embedding), GELU / ELU in
    //          registers; the fp32 tile goes to shared memory (the operand ring is idle once accum_bar fires);
    // phase 2: warp <-> its own 32 rows with the lanes spread over the columns, so residual loads and every
    //          store are coalesced, rows are processed four at a time to keep loads in flight, and the
    //          LayerNorm feeding the next GEMM is a warp reduction.
    const int q = warp & 3, row = q * 32 + lane;
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
    const GemmEpi& e = p.e;
    const int nch = p.n_tile / 16;
    const int pitch = p.n_tile + 4;            // floats: row-per-lane float4 stores stay bank-conflict free
    float* tile = reinterpret_cast<float*>(smem);
    mbar_wait_spin(accum_bar, 0);
    tc_fence_after();
    if (threadIdx.x == 64) GEMM_STAMP(4);
    float mean1 = 0.f, rstd1 = 1.f;
    if (e.ln_g) {
      float s1 = 0.f, s2 = 0.f;
      for (int c = 0; c < nch; ++c) {
        float v[16];
        tmem_ld16(trow + c * 16, v);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float x = v[i] + (e.bias ? __ldg(e.bias + n0 + c * 16 + i) : 0.f);
          s1 += x; s2 += x * x;
        }
      }
      mean1 = s1 / p.n_tile;
      rstd1 = rsqrtf(fmaxf(s2 / p.n_tile - mean1 * mean1, 0.f) + 1e-5f);
    }
    for (int c = 0; c < nch; ++c) {
      float v[16];
      tmem_ld16(trow + c * 16, v);
      const int nb = n0 + c * 16;
      const int bb0 = e.bias_mod > 0 ? nb % e.bias_mod : nb;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        float x = v[i] + (e.bias ? __ldg(e.bias + bb0 + i) : 0.f);
        if (e.ln_g) {
          x = mish_f((x - mean1) * rstd1 * __ldg(e.ln_g + nb + i) + __ldg(e.ln_b + nb + i));
          if (e.temb) x += __ldg(e.temb + (size_t)b * p.N + nb + i);
        }
        if (e.gelu) x = gelu_f(x);
        if (e.elu) x = x > 0.f ? x : expm1f(x);
        v[i] = x;
      }
      float4* dst = reinterpret_cast<float4*>(tile + (size_t)row * pitch + c * 16);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    __syncwarp();
    if (threadIdx.x == 64) GEMM_STAMP(5);
    // ---- phase 2 ----
    const int ngrp = p.n_tile / 4;             // float4 column groups per row; lane owns groups lane and lane + 32
    float g2[2][4], b2[2][4], sa[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int gi = lane + 32 * i;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool ok = gi < ngrp;
        g2[i][k] = (ok && e.out_ln) ? __ldg(e.ln2_g + n0 + gi * 4 + k) : 0.f;
        b2[i][k] = (ok && e.out_ln) ? __ldg(e.ln2_b + n0 + gi * 4 + k) : 0.f;
        sa[i][k] = (ok && e.snake_alpha) ? __ldg(e.snake_alpha + n0 + gi * 4 + k) : 1.f;
      }
    }
    for (int r0 = 0; r0 < 32; r0 += 4) {
      float4 v[4][2];
      bool okr[4];
      size_t grow[4];
#pragma unroll
      for (int rr = 0; rr < 4; ++rr) {
        const int r = q * 32 + r0 + rr, t = t0 + r;
        okr[rr] = t < p.M_out;
        grow[rr] = (size_t)b * p.M_out + (okr[rr] ? t : 0);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int gi = lane + 32 * i;
          v[rr][i] = (gi < ngrp) ? *reinterpret_cast<const float4*>(tile + (size_t)r * pitch + gi * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      if (e.resid) {
        float4 rv[4][2];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int gi = lane + 32 * i;
            rv[rr][i] = (okr[rr] && gi < ngrp) ? *reinterpret_cast<const float4*>(e.resid + grow[rr] * e.ld_resid + n0 + gi * 4)
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) { v[rr][i].x += rv[rr][i].x; v[rr][i].y += rv[rr][i].y; v[rr][i].z += rv[rr][i].z; v[rr][i].w += rv[rr][i].w; }
      }
      if (e.resid2) {
        float4 rv[4][2];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int gi = lane + 32 * i;
            rv[rr][i] = (okr[rr] && gi < ngrp) ? *reinterpret_cast<const float4*>(e.resid2 + grow[rr] * e.ld_resid2 + n0 + gi * 4)
                                              : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
          for (int i = 0; i < 2; ++i) { v[rr][i].x += rv[rr][i].x; v[rr][i].y += rv[rr][i].y; v[rr][i].z += rv[rr][i].z; v[rr][i].w += rv[rr][i].w; }
      }
#pragma unroll
      for (int rr = 0; rr < 4; ++rr) {
        float mean2 = 0.f, rstd2 = 1.f;
        if (e.out_ln) {
          float s1 = 0.f, s2 = 0.f;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            s1 += v[rr][i].x + v[rr][i].y + v[rr][i].z + v[rr][i].w;
            s2 += v[rr][i].x * v[rr][i].x + v[rr][i].y * v[rr][i].y + v[rr][i].z * v[rr][i].z + v[rr][i].w * v[rr][i].w;
          }
          s1 = warp_sum(s1); s2 = warp_sum(s2);
          mean2 = s1 / p.n_tile;
          rstd2 = rsqrtf(fmaxf(s2 / p.n_tile - mean2 * mean2, 0.f) + 1e-5f);
        }
        if (!okr[rr]) continue;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int gi = lane + 32 * i;
          if (gi >= ngrp) continue;
          const int nb = n0 + gi * 4;
          const float4 x = v[rr][i];
          if (e.out_f32) {
            const long long off = (long long)(grow[rr] * e.ld_f32 + nb) + e.out_shift;
            if (e.out_limit == 0 || (off >= 0 && off + 4 <= e.out_limit)) *reinterpret_cast<float4*>(e.out_f32 + off) = x;
          }
          if (e.out_bf16) {
            float y[4] = {x.x, x.y, x.z, x.w};
            if (e.snake_alpha) {
#pragma unroll
              for (int k = 0; k < 4; ++k) { const float sn = sinf(y[k] * sa[i][k]); y[k] = y[k] + (1.0f / (sa[i][k] + 1e-9f)) * sn * sn; }
            } else if (e.lrelu != 0.f) {
#pragma unroll
              for (int k = 0; k < 4; ++k) y[k] = y[k] > 0.f ? y[k] : y[k] * e.lrelu;
            }
            *reinterpret_cast<uint2*>(e.out_bf16 + grow[rr] * e.ld_bf16 + nb) = make_uint2(pack_bf16x2(y[0], y[1]), pack_bf16x2(y[2], y[3]));
          }
          if (e.out_ln) {
            const float y0 = (x.x - mean2) * rstd2 * g2[i][0] + b2[i][0], y1 = (x.y - mean2) * rstd2 * g2[i][1] + b2[i][1];
            const float y2 = (x.z - mean2) * rstd2 * g2[i][2] + b2[i][2], y3 = (x.w - mean2) * rstd2 * g2[i][3] + b2[i][3];
            *reinterpret_cast<uint2*>(e.out_ln + grow[rr] * e.ld_ln + nb) = make_uint2(pack_bf16x2(y0, y1), pack_bf16x2(y2, y3));
          }
        }
      }
    }
    if (threadIdx.x == 64) GEMM_STAMP(6);
    tc_fence_before();
  }
  __syncthreads();
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
  CBX_CHECK(!(epi.ln_g || epi.out_ln) || n_tile == N, "gemm_make: row LayerNorm needs the whole row in one tile (N=%d)", N);
    d->B = B; d->L = L; d->N = N; d->K = K; d->taps = taps; d->n_tile = n_tile; d->epi = epi;
  d->dil = dil; d->pad = pad; d->M_out = M_out;
  if (make_tmap_bf16_3d(&d->tmA, A, K, L, B, lda, (uint64_t)lda * L, BK, BM)) return CBX_ERR;
  if (make_tmap_bf16_3d(&d->tmW, W, (uint64_t)taps * K, N, 1, (uint64_t)taps * K, (uint64_t)taps * K * N, BK, n_tile)) return CBX_ERR;
  return CBX_OK;
}

static unsigned long long* g_gemm_dbg = nullptr;
void gemm_set_debug(unsigned long long* buf) { g_gemm_dbg = buf; }

int gemm_launch(const GemmDesc& d, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    CBX_CUDA(cudaFuncSetAttribute(s3_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
    attr_set = true;
  }
  GemmParams p;
  p.tmA = d.tmA; p.tmW = d.tmW;
  p.B = d.B; p.L = d.L; p.N = d.N; p.K = d.K; p.taps = d.taps; p.n_tile = d.n_tile;
  p.dil = d.dil; p.pad = d.pad; p.M_out = d.M_out;
  p.tiles_l = (d.M_out + BM - 1) / BM;
  p.e = d.epi;
  p.dbg = g_gemm_dbg;
  if (g_gemm_dbg) g_gemm_dbg += 8;
  dim3 grid(d.B * p.tiles_l, d.N / d.n_tile);
  s3_gemm_kernel<<<grid, GEMM_THREADS, GEMM_SMEM, st>>>(p);
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

}  // namespace cbx
