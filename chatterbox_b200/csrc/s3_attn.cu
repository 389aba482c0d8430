// Full (bidirectional) multi-head attention of the S3Gen estimator's BasicTransformerBlock
// (reference: matcha/transformer.py:196-204 -> diffusers 0.29.0 Attention/AttnProcessor2_0: 8 heads x 64,
// scale 1/8, key-padding bias only — all-zero for one un-padded utterance, decoder.py:269-270).
// Flash-attention structure: a CTA owns 64 query rows of one (batch, head); K/V tiles of 64 keys are
// stream through a cp.async pipeline in shared memory; TWO groups of 4 warps share a CTA and take alternate key
// tiles (the per-tile dependency chain, not bandwidth, bounds this kernel: splitting the keys halves it), their
// online-softmax states are merged through shared memory at the end; S = QK^T and O += PV run on mma.sync m16n8k16 bf16
// with fp32 accumulation and an fp32 online softmax.  (Round-1 kernel: the legacy tensor-core path;
// the tcgen05 version is listed in DESIGN.md as the next step for this stage.)
#include "s3_gemm.cuh"
#include <cstdlib>

namespace cbx {
namespace {

constexpr int AT_BM = 64, AT_BN = 64, AT_D = 64, AT_LD = 72;   // padded smem row stride (elements)
constexpr int AT_G = 2;                                         // warp groups (key split inside the CTA)
constexpr int AT_NST = 2 * AT_G;                                // staged K/V tiles: two rounds of AT_G tiles
constexpr int AT_TILE = AT_BN * AT_LD;                          // elements of one staged tile
constexpr int AT_SMEM = (1 + 2 * AT_NST) * AT_TILE * 2;         // Q + NST x (K, V), bytes

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem)), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}

// qkv: bf16 [B][L][ld_qkv] with q at column 0, k at 512, v at 1024 (+ head*64); out: bf16 [B][L][512]
__global__ void __launch_bounds__(128 * AT_G) s3_attn_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                                                      int L, int ld_qkv, const int* __restrict__ lens) {
  extern __shared__ __align__(16) __nv_bfloat16 at_smem[];
  __nv_bfloat16* sQ = at_smem;
  __nv_bfloat16* sKb = at_smem + AT_TILE;                 // [NST][tile]
  __nv_bfloat16* sVb = at_smem + (1 + AT_NST) * AT_TILE;  // [NST][tile]
  const int tid = threadIdx.x, grp = tid >> 7, warp = (tid >> 5) & 3, lane = tid & 31;
  const int q0 = blockIdx.x * AT_BM, head = blockIdx.y, b = blockIdx.z;
  const __nv_bfloat16* base = qkv + (size_t)b * L * ld_qkv + head * AT_D;

  auto load_tile = [&](__nv_bfloat16* dst, int col_off, int row0) {
#pragma unroll
    for (int i = 0; i < 4 / AT_G; ++i) {
      const int ch = tid + i * 128 * AT_G, r = ch >> 3, c8 = ch & 7;
      const bool ok = row0 + r < L;
      cp_async16(dst + r * AT_LD + c8 * 8, base + (size_t)(ok ? row0 + r : 0) * ld_qkv + col_off + c8 * 8, ok);
    }
  };
  pdl_launch_dependents();
  pdl_wait();                 // q|k|v come from the GEMM just before
  // key-padding mask of a padded batch (decoder.py:25-33, utils/mask.py:167-193): row b attends to its first Lk keys only;
  // key tiles past Lk are not even loaded.  lens == NULL: every row is L long (one un-padded utterance).
  const int Lk = lens ? min(max(lens[b], 1), L) : L;
  const int ntiles = (Lk + AT_BN - 1) / AT_BN;
  load_tile(sQ, 0, q0);
  const int nrounds = (ntiles + AT_G - 1) / AT_G;        // a round = AT_G consecutive key tiles, one per warp group
#pragma unroll
  for (int t = 0; t < AT_G; ++t) {
    if (t < ntiles) {
      load_tile(sKb + t * AT_TILE, 512, t * AT_BN);
      load_tile(sVb + t * AT_TILE, 1024, t * AT_BN);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");        // one commit group per round

  float o[8][4], m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  uint32_t qf[4][4];
  const float sc = 0.125f * 1.4426950408889634f;   // 1/sqrt(64) * log2(e)

  for (int rd = 0; rd < nrounds; ++rd) {
    const int kt = rd * AT_G + grp;            // this group's key tile of the round
    const __nv_bfloat16* sK = sKb + (kt % AT_NST) * AT_TILE;
    const __nv_bfloat16* sV = sVb + (kt % AT_NST) * AT_TILE;
#pragma unroll
    for (int t = 0; t < AT_G; ++t) {           // next round's tiles into the other half of the ring (freed by the barrier
      const int nt = (rd + 1) * AT_G + t;      // that ended round rd - 1)
      if (nt < ntiles) {
        load_tile(sKb + (nt % AT_NST) * AT_TILE, 512, nt * AT_BN);
        load_tile(sVb + (nt % AT_NST) * AT_TILE, 1024, nt * AT_BN);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");      // round rd (and Q) have landed
    __syncthreads();
    if (rd == 0) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks)
        ldsm_x4(qf[ks], sQ + (warp * 16 + (lane & 15)) * AT_LD + ks * 16 + (lane >> 4) * 8);
    }
    if (kt < ntiles) {
    // ---- S = Q K^T (16 x 64 per warp) ----
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {     // pairs of 8-key tiles
        uint32_t kb[4];
        ldsm_x4(kb, sK + ((np * 2 + (lane >> 4)) * 8 + (lane & 7)) * AT_LD + ks * 16 + ((lane >> 3) & 1) * 8);
        mma_bf16_16816(s[np * 2], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], kb[0], kb[1]);
        mma_bf16_16816(s[np * 2 + 1], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], kb[2], kb[3]);
      }
    }
    // ---- scale, tail mask, online softmax (rows g = lane>>2 and g + 8) ----
    const int kbase = kt * AT_BN + (lane & 3) * 2;
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int key = kbase + i * 8 + (e & 1);
        const float v = (key < Lk) ? s[i][e] * sc : -INFINITY;
        s[i][e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
    }
    float corr[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float mn = fmaxf(m_run[r], mx[r]);
      corr[r] = exp2f(m_run[r] - mn);
      m_run[r] = mn;
    }
    float ls[2] = {0.f, 0.f};
    uint32_t pf[4][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float p0 = exp2f(s[i][0] - m_run[0]), p1 = exp2f(s[i][1] - m_run[0]);
      const float p2 = exp2f(s[i][2] - m_run[1]), p3 = exp2f(s[i][3] - m_run[1]);
      ls[0] += p0 + p1; ls[1] += p2 + p3;
      pf[i >> 1][(i & 1) * 2] = pack_bf16x2(p0, p1);
      pf[i >> 1][(i & 1) * 2 + 1] = pack_bf16x2(p2, p3);
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) l_run[r] = l_run[r] * corr[r] + ls[r];
#pragma unroll
    for (int i = 0; i < 8; ++i) { o[i][0] *= corr[0]; o[i][1] *= corr[0]; o[i][2] *= corr[1]; o[i][3] *= corr[1]; }
    // ---- O += P V ----
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {     // pairs of 8-wide d tiles
        uint32_t vb[4];
        ldsm_x4_t(vb, sV + (kk * 16 + (lane & 15)) * AT_LD + (dp * 2 + (lane >> 4)) * 8);
        mma_bf16_16816(o[dp * 2], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], vb[0], vb[1]);
        mma_bf16_16816(o[dp * 2 + 1], pf[kk][0], pf[kk][1], pf[kk][2], pf[kk][3], vb[2], vb[3]);
      }
    }
    }
    __syncthreads();
  }
  // ---- merge the warp groups' partial states (group g > 0 -> shared memory -> group 0) ----
  {
    float* xch = reinterpret_cast<float*>(sKb);                 // ring is idle now: [128 threads][36 floats]
    const int t128 = tid & 127;
    if (grp == 1) {
      float* d = xch + t128 * 36;
#pragma unroll
      for (int i = 0; i < 8; ++i) { d[i * 4] = o[i][0]; d[i * 4 + 1] = o[i][1]; d[i * 4 + 2] = o[i][2]; d[i * 4 + 3] = o[i][3]; }
      d[32] = m_run[0]; d[33] = m_run[1]; d[34] = l_run[0]; d[35] = l_run[1];
    }
    __syncthreads();
    if (grp != 0) return;
    const float* d = xch + t128 * 36;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const float m1 = d[32 + r], l1 = d[34 + r];
      const float mn = fmaxf(m_run[r], m1);
      const float c0 = (m_run[r] == -INFINITY) ? 0.f : exp2f(m_run[r] - mn), c1 = (m1 == -INFINITY) ? 0.f : exp2f(m1 - mn);
      l_run[r] = l_run[r] * c0 + l1 * c1;
      m_run[r] = mn;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        o[i][r * 2] = o[i][r * 2] * c0 + d[i * 4 + r * 2] * c1;
        o[i][r * 2 + 1] = o[i][r * 2 + 1] * c0 + d[i * 4 + r * 2 + 1] * c1;
      }
    }
  }
  // ---- normalise and store ----
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const int g = lane >> 2, c2 = (lane & 3) * 2;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int t = q0 + warp * 16 + g + r * 8;
    if (t < L) {
      const float inv = 1.0f / l_run[r];
      __nv_bfloat16* dst = out + ((size_t)b * L + t) * 512 + head * AT_D + c2;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        *reinterpret_cast<uint32_t*>(dst + i * 8) = pack_bf16x2(o[i][r * 2] * inv, o[i][r * 2 + 1] * inv);
    }
  }
}

}  // namespace

int s3_attn_tc_launch(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st);
int s3_attn_launch_mma(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st);

int s3_attn_launch(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st) {
  // default: the tcgen05 kernel (csrc/s3_attn_tc.cu); CBX_ATTN_TC=0 keeps the mma.sync kernel below for A/B
  static const bool use_tc = [] { const char* v = getenv("CBX_ATTN_TC"); return !(v && v[0] == '0'); }();
  if (use_tc) return s3_attn_tc_launch(qkv, out, B, L, ld_qkv, lens, st);
  return s3_attn_launch_mma(qkv, out, B, L, ld_qkv, lens, st);
}

int s3_attn_launch_mma(const __nv_bfloat16* qkv, __nv_bfloat16* out, int B, int L, int ld_qkv, const int* lens, cudaStream_t st) {
  dim3 grid((L + AT_BM - 1) / AT_BM, 8, B);
  static CbxPerDeviceOnce once;
  if (once.first())
    CBX_CUDA(cudaFuncSetAttribute(s3_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = dim3(128 * AT_G); cfg.dynamicSmemBytes = AT_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_attn_kernel, qkv, out, L, ld_qkv, lens));
  CBX_CUDA(cudaGetLastError());
  return CBX_OK;
}

}  // namespace cbx
