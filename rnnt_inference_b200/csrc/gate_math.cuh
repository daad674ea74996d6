// gate_math.cuh — the reference's fp16 / fp32 gate arithmetic, restated operation by operation for CUDA.
//
// fp16 path (encoder):  _mm512_sigmoid_ph / _mm512_tanh_ph (tpps/el_common_intrin.hpp:95-152),
// _mm512_scale_min128max_i8_ph (:295-303), lstm_postop_tpp::ref (tpps/lstm_postop_tpp.cpp:12-98).
// Each AVX512-FP16 instruction is one IEEE binary16 operation, so it maps 1:1 onto __hmul2_rn / __hadd2_rn /
// __hfma2 (the _rn forms also forbid the compiler from contracting mul+add into fma, which would change
// results).  VRCPPH is not correctly rounded; it is reproduced from the 1024-entry mantissa table
// include/rnnt_rcp_ph_lut.h (binade-invariant, dumped from hardware).
// fp32 path (prediction net): _mm512_sigmoid_ps / _mm512_tanh_ps (:66-93,154-185) with VRCP14PS replaced by
// RN(1/x) — the deviation the oracle declares.
#pragma once
#include "common.cuh"

namespace rnnt {

__device__ __forceinline__ uint32_t h2_bits(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ __half2 bits_h2(uint32_t b) { return *reinterpret_cast<__half2*>(&b); }

// VRCPPH on both halves; lut = 1024 x uint16 in shared memory.  Inputs are positive normal numbers
// (denominators >= 495) or +inf (result then discarded by the |x| > 4.97 clamp).
__device__ __forceinline__ __half2 rcp_ph2(__half2 d, const uint16_t* lut) {
  const uint32_t b = h2_bits(d);
  const uint32_t lo = b & 0xffffu, hi = b >> 16;
  const uint32_t rlo = (uint32_t)lut[lo & 0x3ffu] + ((15u - ((lo >> 10) & 31u)) << 10);
  const uint32_t rhi = (uint32_t)lut[hi & 0x3ffu] + ((15u - ((hi >> 10) & 31u)) << 10);
  return bits_h2((rlo & 0xffffu) | (rhi << 16));
}

struct HalfConsts {
  __half2 c495, c60, c225, c10, c1_21, hi, lo, one, neg_one, half;
};
__device__ __forceinline__ HalfConsts make_half_consts() {
  HalfConsts k;
  k.c495 = __float2half2_rn(495.0f);
  k.c60 = __float2half2_rn(60.0f);
  k.c225 = __float2half2_rn(225.0f);
  k.c10 = __float2half2_rn(10.0f);
  k.c1_21 = __float2half2_rn((float)(1.0 / 21.0));
  k.hi = __float2half2_rn(4.97f);
  k.lo = __float2half2_rn(-4.97f);
  k.one = __float2half2_rn(1.0f);
  k.neg_one = __float2half2_rn(-1.0f);
  k.half = __float2half2_rn(0.5f);
  return k;
}

// rational core shared by tanh_ph(x) and sigmoid_ph(x) (which evaluates it at x/2)
__device__ __forceinline__ __half2 pade_clamped(__half2 x, const HalfConsts& k, const uint16_t* lut) {
  const __half2 x2 = __hmul2_rn(x, x);
  const __half2 t1n = __hadd2_rn(x2, k.c60);
  const __half2 t2n = __hfma2(t1n, x2, k.c495);
  const __half2 nom = __hmul2_rn(t2n, x);
  const __half2 t1d = __hfma2(k.c1_21, x2, k.c10);
  const __half2 t2d = __hfma2(t1d, x2, k.c225);
  const __half2 den = __hfma2(t2d, x2, k.c495);
  __half2 r = __hmul2_rn(nom, rcp_ph2(den, lut));
  const uint32_t mh = __hgt2_mask(x, k.hi);   // ordered >, 0xffff per true half
  const uint32_t ml = __hlt2_mask(x, k.lo);
  uint32_t rb = h2_bits(r);
  rb = (rb & ~mh) | (h2_bits(k.one) & mh);
  rb = (rb & ~ml) | (h2_bits(k.neg_one) & ml);
  return bits_h2(rb);
}
__device__ __forceinline__ __half2 tanh_ph2(__half2 x, const HalfConsts& k, const uint16_t* lut) {
  return pade_clamped(x, k, lut);
}
__device__ __forceinline__ __half2 sigmoid_ph2(__half2 x, const HalfConsts& k, const uint16_t* lut) {
  const __half2 xh = __hmul2_rn(k.half, x);
  __half2 r = pade_clamped(xh, k, lut);
  r = __hadd2_rn(r, k.one);
  return __hmul2_rn(r, k.half);
}

// Whole-function table of pade_clamped: it is an odd function of x in exact binary16 arithmetic (x only enters through
// x*x and one final multiply; the clamp thresholds are symmetric), so tab[bits(|x|)] with the sign bit OR-ed back gives
// the identical result for every input with |x| bits <= 0x44F9; larger magnitudes (incl. inf) clamp to entry 0x44F9 =
// 1.0.  The table is filled by the kernel itself with pade_clamped, i.e. it is exact by construction.  NaN inputs (never
// produced by finite accumulators) would map to +-1 instead of NaN.
__device__ __forceinline__ __half2 tanh_tab2(__half2 x, uint32_t tab_smem) {
  const uint32_t b = h2_bits(x);
  uint32_t idx = __vminu2(b & 0x7fff7fffu, 0x44f944f9u);
  idx += idx;                                   // byte offsets of both halves at once (each < 0x8000: no carry across)
  uint32_t lo, hi;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(lo) : "r"(tab_smem + (idx & 0xffffu)));
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(hi) : "r"(tab_smem + (idx >> 16)));
  return bits_h2(__byte_perm(lo, hi, 0x5410) | (b & 0x80008000u));
}
__device__ __forceinline__ __half2 sigmoid_tab2(__half2 x, const HalfConsts& k, uint32_t tab_smem) {
  __half2 r = tanh_tab2(__hmul2_rn(k.half, x), tab_smem);
  r = __hadd2_rn(r, k.one);
  return __hmul2_rn(r, k.half);
}

// _mm512_scale_min128max_i8_ph: fp16 multiply by the fp16-rounded scale, round to nearest-even integer,
// min(127), max(-128); returns the two int8 values in the low 16 bits (byte0 = .x, byte1 = .y).
// The clamp bounds are integers and rint is monotonic, so clamping first gives the same result; the clamped value v
// in [-128, 127] is then rounded by ONE fp16 addition: v + 1536 lies in [1024, 2048) where the spacing is 1, the add
// rounds to nearest-even exactly like rint (1536 is even, so ties land on the same integers), and the low byte of
// the sum's mantissa (1536 - 1024 = 512 = 0 mod 256) is the two's-complement int8.
__device__ __forceinline__ uint32_t quant_ph2(__half2 h, __half2 scale) {
  __half2 m = __hmul2_rn(h, scale);
  m = __hmin2(m, __float2half2_rn(127.f));
  m = __hmax2(m, __float2half2_rn(-128.f));
  const uint32_t r = h2_bits(__hadd2_rn(m, __float2half2_rn(1536.f)));
  return __byte_perm(r, 0u, 0x4420);
}

// dequant_float_out_accum (tpps/tile_product_tpp.hpp:308-345): int32 add FIRST, then cvt, + bias, * scale —
// two separately rounded fp32 operations (no fma), then VCVTPS2PH (RNE).
__device__ __forceinline__ float dequant_gate(int acc_hh, int acc_ih, float bias, float rb_scale) {
  return __fmul_rn(__fadd_rn(__int2float_rn(acc_hh + acc_ih), bias), rb_scale);
}

// ---------------------------------------------------------------- fp32 path (prediction network)
__device__ __forceinline__ float pade_ps_clamped(float x) {
  const float c0 = 4826.25f, c1 = 618.75f, c2 = 13.5f, c3 = (float)(1.0 / 28.0), c1d = 2227.5f, c2d = 112.5f;
  const float x2 = __fmul_rn(x, x);
  const float t1n = __fmaf_rn(c3, x2, c2);
  const float t2n = __fmaf_rn(t1n, x2, c1);
  const float t3n = __fmaf_rn(t2n, x2, c0);
  const float nom = __fmul_rn(t3n, x);
  const float t1d = __fadd_rn(x2, c2d);
  const float t2d = __fmaf_rn(t1d, x2, c1d);
  const float den = __fmaf_rn(t2d, x2, c0);
  float r = __fmul_rn(nom, __frcp_rn(den));
  if (x > 4.97f) r = 1.0f;
  if (x < -4.97f) r = -1.0f;
  return r;
}
__device__ __forceinline__ float tanh_ps(float x) { return pade_ps_clamped(x); }
__device__ __forceinline__ float sigmoid_ps(float x) {
  float r = pade_ps_clamped(__fmul_rn(0.5f, x));
  r = __fadd_rn(r, 1.0f);
  return __fmul_rn(r, 0.5f);
}

__device__ __forceinline__ float bf16_bits_to_float(uint32_t b) { return __uint_as_float(b << 16); }
__device__ __forceinline__ uint32_t float_to_bf16_bits(float f) {
  return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(f));
}

}  // namespace rnnt
