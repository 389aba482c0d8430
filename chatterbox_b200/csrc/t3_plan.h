// Static work plan of the T3 decode megakernel: which 16 KB weight "units" each CTA consumes, in
// which order.  One definition shared by the device kernel and (through the cbx_t3_plan_* exports)
// by the host-side weight packer, so the per-CTA weight streams in HBM are laid out in exactly the
// order the TMA producer walks them.
//
// A unit is 16 weight rows x 1024 k (bf16) = 32 KB.  Per decoder layer (HF LlamaDecoderLayer as
// configured by reference src/chatterbox/models/t3/llama_configs.py:1-33):
//   PH_QKV  : 3072 x 1024 fused q|k|v          -> 192 units, unit u = rows [16u, 16u+16)
//   PH_O    : 1024 x 1024 o_proj               -> 64 units
//   PH_GU   : 8192 x 1024 gate/up interleaved  -> 512 units, unit u = gate rows 8u..8u+7 then up rows 8u..8u+7
//   PH_DOWN : 1024 x 4096 down_proj            -> 256 units, unit u = r*4+s: rows [16r,16r+16), k-slice [1024s, 1024s+1024)
// and once per step
//   PH_HEAD : 8194(->8208) x 1024 speech_head  -> 513 units (reference t3/t3.py:64)
// Units of a phase are dealt round-robin over the G CTAs starting at a per-phase rotation chosen so
// that the per-layer totals differ by at most one unit between CTAs.  In PH_DOWN all units of a CTA
// share one k-slice so its activation fragments stay in registers.
#pragma once

#ifdef __CUDACC__
#define CBX_HD __host__ __device__ __forceinline__
#else
#define CBX_HD static inline
#endif

enum { PH_QKV = 0, PH_O = 1, PH_GU = 2, PH_DOWN = 3, PH_HEAD = 4, PH_COUNT = 5 };

#define T3_H 1024
#define T3_NH 16
#define T3_HD 64
#define T3_INTER 4096
#define T3_VOCAB 8194
#define T3_VOCAB_PAD 8208  // speech_head rows padded to whole units
#define T3_UNIT_ROWS 16
#define T3_UNIT_BYTES 32768
#define T3_MAXU 4          // max units of one phase on one CTA (needs G >= 129)
#define T3_KW 8            // warps of a CTA; they split the 1024-wide k range of a unit (the fragment layout depends on it)
#define T3_DOWN_ROWG 64    // 1024 / 16 row groups of down_proj
#define T3_DOWN_KS 4       // k-slices of down_proj

CBX_HD int t3_phase_units(int ph) {
  return ph == PH_QKV ? 192 : ph == PH_O ? 64 : ph == PH_GU ? 512 : ph == PH_DOWN ? 256 : 513;
}
CBX_HD int t3_phase_rot(int ph, int G) {
  // cumulative unit count of the preceding phases of the layer, mod G
  return ph == PH_QKV ? 0 : ph == PH_O ? (192 % G) : ph == PH_GU ? ((192 + 64) % G)
       : ph == PH_DOWN ? ((192 + 64 + 512) % G) : 0;
}
// number of units of phase `ph` that CTA `c` of `G` consumes
CBX_HD int t3_plan_count(int ph, int c, int G) {
  int ce = c - t3_phase_rot(ph, G);
  if (ce < 0) ce += G;
  if (ph == PH_DOWN) {
    int n4 = G / T3_DOWN_KS;  // CTAs per k-slice
    if (ce >= n4 * T3_DOWN_KS) return 0;
    int rbase = ce / T3_DOWN_KS;
    return rbase < T3_DOWN_ROWG ? (T3_DOWN_ROWG - 1 - rbase) / n4 + 1 : 0;
  }
  int U = t3_phase_units(ph);
  return ce < U ? (U - 1 - ce) / G + 1 : 0;
}
// j-th unit id of phase `ph` on CTA `c`
CBX_HD int t3_plan_unit(int ph, int c, int G, int j) {
  int ce = c - t3_phase_rot(ph, G);
  if (ce < 0) ce += G;
  if (ph == PH_DOWN) {
    int n4 = G / T3_DOWN_KS;
    int s = ce % T3_DOWN_KS, r = ce / T3_DOWN_KS + j * n4;
    return r * T3_DOWN_KS + s;
  }
  return ce + j * G;
}
CBX_HD int t3_plan_units_per_layer(int c, int G) {
  return t3_plan_count(PH_QKV, c, G) + t3_plan_count(PH_O, c, G) + t3_plan_count(PH_GU, c, G) +
         t3_plan_count(PH_DOWN, c, G);
}
