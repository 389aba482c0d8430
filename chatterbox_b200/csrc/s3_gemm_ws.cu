// Weight-stationary persistent tcgen05 GEMM for the padded-batch estimator (reference: matcha/transformer.py:276-316 — the
// q|k|v projection and the feed-forward up-projection + GELU of BasicTransformerBlock at M = 2 nb L ~ 10^5 rows).
//
// At that size the one-tile-per-CTA kernel of s3_gemm.cu is bound by what it does NOT overlap: every CTA loads its own copy
// of the 128 KB weight tile, computes for ~1 us and then spends 4-5 us draining and storing, one CTA per SM (8 us per tile,
// ~280 TFLOP/s).  Here a CTA keeps ONE column tile's weights [256][256] in shared memory for its whole life and walks the
// row tiles of that column:
//   warp 0      TMA producer: the weight tile once (before the programmatic-launch wait), then the A k-blocks [128][64] of
//               one row tile after the other through a 4-stage ring;
//   warp 1      MMA issuer: M128 N256 K16 x 16 per row tile into one of TWO TMEM accumulators (2 x 256 columns);
//   warps 2-17  epilogue: thread <-> accumulator row, 64 columns per warp (four warps per TMEM lane quarter): tcgen05.ld,
//               bias, GELU, bf16, through a per-warp staging tile so that global stores are whole 128-byte lines — while the
//               MMA warp is already filling the other accumulator with the next row tile.
// K = 256 and a 256-column tile only (the two shapes above); everything else stays on s3_gemm.cu.  CBX_GEMM_WS=0 disables it.
#include "s3_gemm.cuh"
#include <cstdlib>

namespace cbx {
namespace {

constexpr int WM = 128, WN = 256, WK = 256, WBK = 64, WKB = WK / WBK;
constexpr int W_KB = WN * WBK * 2;            // 32 KB: [256 rows][64 k]
constexpr int A_KB = WM * WBK * 2;            // 16 KB
constexpr int WSTG = 4;                       // A ring stages (one row tile)
constexpr int WS_OFF_A = WKB * W_KB;          // 128 KB of weights, then the ring
constexpr int WS_EPI_WARPS = 16;
constexpr int WS_OFF_STG = WS_OFF_A + WSTG * A_KB;       // 2 KB per epilogue warp: [16 rows][64 bf16], 128-byte rows, swizzled
constexpr int WS_OFF_BAR = WS_OFF_STG + WS_EPI_WARPS * 2048;
constexpr int WS_SMEM = WS_OFF_BAR + 256 + WN * 4 + 1024;
static_assert(WS_SMEM <= 232448, "shared memory per CTA");
constexpr int WS_THREADS = 32 * (2 + WS_EPI_WARPS);

struct WsParams {
  CUtensorMap tmA, tmW;
  int tiles_l, n_row_tiles, ncols, per_col, M_out;
  const float* bias;
  int gelu;
  __nv_bfloat16* out;
  int ld_out;
};

__global__ void __launch_bounds__(WS_THREADS, 1) s3_gemm_ws_kernel(const __grid_constant__ WsParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sW = smem;
  uint8_t* sA = smem + WS_OFF_A;
  uint64_t* bar_w = reinterpret_cast<uint64_t*>(smem + WS_OFF_BAR);
  uint64_t* full = bar_w + 1;            // [WSTG]
  uint64_t* empty = full + WSTG;         // [WSTG]
  uint64_t* acc_full = empty + WSTG;     // [2]
  uint64_t* acc_empty = acc_full + 2;    // [2] one arrive per epilogue warp
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  float* sBias = reinterpret_cast<float*>(smem + WS_OFF_BAR + 256);       // [256] this column tile's bias

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = blockIdx.x % p.ncols, first = blockIdx.x / p.ncols;
  const int n0 = col * WN;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    for (int i = 0; i < WSTG; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], WS_EPI_WARPS); }
    mbar_fence_init();
    tma_prefetch_desc(&p.tmA); tma_prefetch_desc(&p.tmW);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  if (threadIdx.x >= 64 && threadIdx.x < 64 + WN) sBias[threadIdx.x - 64] = p.bias ? __ldg(p.bias + n0 + threadIdx.x - 64) : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_arrive_expect_tx(bar_w, WKB * W_KB);
      for (int kb = 0; kb < WKB; ++kb) tma_load_3d(sW + kb * W_KB, &p.tmW, bar_w, kb * WBK, n0, 0);
      pdl_wait();                        // the activations come from the kernel before
      int s = 0, ph = 0;
      for (int rt = first; rt < p.n_row_tiles; rt += p.per_col) {
        const int b = rt / p.tiles_l, t0 = (rt % p.tiles_l) * WM;
        for (int kb = 0; kb < WKB; ++kb) {
          mbar_wait_spin(&empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&full[s], A_KB);
          tma_load_3d(sA + s * A_KB, &p.tmA, &full[s], kb * WBK, t0, b);
          if (++s == WSTG) { s = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(WM, WN);
      mbar_wait_spin(bar_w, 0);
      int s = 0, ph = 0, buf = 0, aph = 0;
      for (int rt = first; rt < p.n_row_tiles; rt += p.per_col) {
        mbar_wait_spin(&acc_empty[buf], aph ^ 1);          // the epilogue has drained this accumulator (first use: free)
        tc_fence_after();
        for (int kb = 0; kb < WKB; ++kb) {
          mbar_wait_spin(&full[s], ph);
          tc_fence_after();
          const uint64_t da = umma_desc_k128(smem_u32(sA + s * A_KB)), db = umma_desc_k128(smem_u32(sW + kb * W_KB));
#pragma unroll
          for (int k = 0; k < WBK / 16; ++k) umma_bf16(tmem_base + buf * WN, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          umma_commit(&empty[s]);
          if (++s == WSTG) { s = 0; ph ^= 1; }
        }
        umma_commit(&acc_full[buf]);
        if (++buf == 2) { buf = 0; aph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue warps: quarter = warp % 4 (the TMEM lanes a warp may touch), 64 columns per warp ----
    const int q = warp & 3, cc = (warp - 2) >> 2;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const float* bias = sBias + cc * 64;
    uint8_t* stg = smem + WS_OFF_STG + (warp - 2) * 2048;
    int buf = 0, aph = 0;
    for (int rt = first; rt < p.n_row_tiles; rt += p.per_col) {
      const int b = rt / p.tiles_l, tq0 = (rt % p.tiles_l) * WM + q * 32;      // first row of this warp's TMEM lane quarter
      mbar_wait_spin(&acc_full[buf], aph);
      tc_fence_after();
      uint32_t w[32];                       // this thread's row, 64 columns, packed bf16
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float v[32];
        tmem_ld32(tmem_base + lane_off + buf * WN + cc * 64 + h * 32, v);
        if (h == 1) {                     // both halves are in registers: the accumulator may be overwritten
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float a = v[2 * i] + bias[h * 32 + 2 * i], c = v[2 * i + 1] + bias[h * 32 + 2 * i + 1];
          if (p.gelu) { a = gelu_f(a); c = gelu_f(c); }
          w[h * 16 + i] = pack_bf16x2(a, c);
        }
      }
      // A thread owns a ROW (TMEM lane): stored straight from registers, a warp instruction would touch 32 different
      // 128-byte lines.  Through a 2 KB staging tile instead, 16 rows at a time: 8 lanes then write one row's 128 contiguous
      // bytes, 4 full lines per instruction.
      __nv_bfloat16* dstq = p.out + ((size_t)b * p.M_out + tq0) * p.ld_out + n0 + cc * 64;
#pragma unroll
      for (int ps = 0; ps < 2; ++ps) {
        if ((lane >> 4) == ps) {
          const int rl = lane & 15;
#pragma unroll
          for (int c = 0; c < 8; ++c)
            *reinterpret_cast<uint4*>(stg + rl * 128 + ((c ^ (rl & 7)) << 4)) = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
        }
        __syncwarp();
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rl = it * 4 + (lane >> 3), c = lane & 7, r = ps * 16 + rl;
          const uint4 x = *reinterpret_cast<const uint4*>(stg + rl * 128 + ((c ^ (rl & 7)) << 4));
          if (tq0 + r < p.M_out) *reinterpret_cast<uint4*>(dstq + (size_t)r * p.ld_out + c * 8) = x;
        }
        __syncwarp();
      }
      if (++buf == 2) { buf = 0; aph ^= 1; }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

// true if the descriptor is one of the two shapes this kernel covers AND large enough to pay off
bool gemm_ws_eligible(const GemmDesc& d) {
  static const bool on = [] { const char* v = getenv("CBX_GEMM_WS"); return !(v && v[0] == '0'); }();
  if (!on || d.hilo || d.taps != 1 || d.K != WK || d.n_tile != WN || d.ksplit != 1 || d.nsplit != 1 || d.M_out != d.L) return false;
  const GemmEpi& e = d.epi;
  if (!e.out_bf16 || e.out_f32 || e.ln_g || e.out_ln || e.temb || e.elu || e.resid || e.resid2 || e.snake_alpha || e.lrelu != 0.f ||
      e.out_hilo || e.swiglu || e.out_shift || e.out_limit || e.bias_mod)
    return false;
  const long long tiles = (long long)d.B * ((d.M_out + WM - 1) / WM) * (d.N / WN);
  return tiles >= 4 * 148;            // fewer tiles: the one-tile-per-CTA kernel fills the GPU in a few waves anyway
}

int gemm_ws_launch(const GemmDesc& d, cudaStream_t st) {
  static CbxPerDeviceOnce once;
  if (once.first()) CBX_CUDA(cudaFuncSetAttribute(s3_gemm_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WS_SMEM));
  WsParams p;
  p.tmA = d.tmA; p.tmW = d.tmW;
  p.tiles_l = (d.M_out + WM - 1) / WM; p.n_row_tiles = d.B * p.tiles_l; p.ncols = d.N / WN; p.M_out = d.M_out;
  p.per_col = 148 / p.ncols;
  if (p.per_col > p.n_row_tiles) p.per_col = p.n_row_tiles;
  p.bias = d.epi.bias; p.gelu = d.epi.gelu; p.out = d.epi.out_bf16; p.ld_out = d.epi.ld_bf16;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(p.ncols * p.per_col); cfg.blockDim = dim3(WS_THREADS); cfg.dynamicSmemBytes = WS_SMEM; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  CBX_CUDA(cudaLaunchKernelEx(&cfg, s3_gemm_ws_kernel, p));
  return CBX_OK;
}

}  // namespace cbx
