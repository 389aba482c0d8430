// Row-wise helpers shared by the GEMM paths of T3 (prompt prefill, csrc/t3_prefill.cu, and the batched decode engine,
// csrc/t3_batch.cu): the Llama RMSNorm written as hi / lo bf16 planes, and the KV slab element types.
#pragma once
#include "common.cuh"

namespace cbx {
namespace t3rows {

constexpr int H = 1024, NH = 16, HD = 64, INTER = 4096, VOCAB = 8194;
constexpr float RMS_EPS = 1e-5f;

// one row per CTA (256 threads x 4 values): out planes [2][rows][1024] = hi / lo of gain * x * rsqrt(mean(x^2) + eps)
static __global__ void __launch_bounds__(256) k_rms_hilo(const float* __restrict__ x, const float* __restrict__ gain,
                                                  __nv_bfloat16* __restrict__ out, long long plane) {
  __shared__ float red[8];
  const int row = blockIdx.x, t = threadIdx.x;
  // programmatic dependent launch (the batched engine launches this kernel with the attribute): let the next kernel's
  // prologue start, and wait for the producer of x before the first read.  Both are no-ops in a plain launch.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const float4 v = *reinterpret_cast<const float4*>(x + (size_t)row * H + t * 4);
  float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  s = warp_sum(s);
  if ((t & 31) == 0) red[t >> 5] = s;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) tot += red[w];
  const float sc = rsqrtf(tot * (1.0f / H) + RMS_EPS);
  const float4 g = __ldg(reinterpret_cast<const float4*>(gain + t * 4));
  uint32_t h0, l0, h1, l1;
  split_hi_lo(g.x * (v.x * sc), g.y * (v.y * sc), h0, l0);
  split_hi_lo(g.z * (v.z * sc), g.w * (v.w * sc), h1, l1);
  __nv_bfloat16* d = out + (size_t)row * H + t * 4;
  *reinterpret_cast<uint2*>(d) = make_uint2(h0, h1);
  *reinterpret_cast<uint2*>(d + plane) = make_uint2(l0, l1);
}

template <typename KVT> struct KVS;
template <> struct KVS<__nv_bfloat16> {
  static __device__ __forceinline__ float round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};
template <> struct KVS<float> {
  static __device__ __forceinline__ float round(float v) { return v; }
  static __device__ __forceinline__ void store(float* p, float v) { *p = v; }
};


}  // namespace t3rows
}  // namespace cbx
