// gen_postop_golden.cpp — produces golden vectors for the fp16 gate math on REAL AVX512-FP16 hardware.
// Test infrastructure.  It is an independent statement of the operation sequence of
// mlperf_plugins/csrc/tpps/lstm_postop_tpp.cpp:12-98 and el_common_intrin.hpp:95-152,295-303 written
// with the same instruction classes (VMULPH/VADDPH/VFMADD*PH/VRCPPH/VCVTPS2PH/VRNDSCALEPH), so the
// scalar oracle's "one IEEE rounding per op + VRCPPH table" model is checked against the silicon.
// Build/run: see tests/golden/make_golden.sh.  Output: raw little-endian arrays on stdout files.
#include <immintrin.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <random>

static inline __m512h tanh_ph(__m512h x) {
  auto c0 = _mm512_set1_ph((_Float16)495.0f), c1 = _mm512_set1_ph((_Float16)60.0f), c1d = _mm512_set1_ph((_Float16)225.0f);
  auto c2d = _mm512_set1_ph((_Float16)10.0f), c3d = _mm512_set1_ph((_Float16)(float)(1.0 / 21.0));
  auto hi = _mm512_set1_ph((_Float16)4.97f), lo = _mm512_set1_ph((_Float16)-4.97f);
  auto x2 = _mm512_mul_ph(x, x);
  auto t1n = _mm512_add_ph(x2, c1);
  auto t2n = _mm512_fmadd_ph(t1n, x2, c0);
  auto nom = _mm512_mul_ph(t2n, x);
  auto t1d = _mm512_fmadd_ph(c3d, x2, c2d);
  auto t2d = _mm512_fmadd_ph(t1d, x2, c1d);
  auto den = _mm512_fmadd_ph(t2d, x2, c0);
  auto rcp = _mm512_rcp_ph(den);
  auto mh = _mm512_cmp_ph_mask(x, hi, _CMP_GT_OQ), ml = _mm512_cmp_ph_mask(x, lo, _CMP_LT_OQ);
  auto r = _mm512_mul_ph(nom, rcp);
  r = _mm512_mask_blend_ph(mh, r, _mm512_set1_ph((_Float16)1.0f));
  r = _mm512_mask_blend_ph(ml, r, _mm512_set1_ph((_Float16)-1.0f));
  return r;
}
static inline __m512h sigmoid_ph(__m512h x) {
  auto half = _mm512_set1_ph((_Float16)0.5f), one = _mm512_set1_ph((_Float16)1.0f);
  auto c0 = _mm512_set1_ph((_Float16)495.0f), c1 = _mm512_set1_ph((_Float16)60.0f), c1d = _mm512_set1_ph((_Float16)225.0f);
  auto c2d = _mm512_set1_ph((_Float16)10.0f), c3d = _mm512_set1_ph((_Float16)(float)(1.0 / 21.0));
  auto hi = _mm512_set1_ph((_Float16)4.97f), lo = _mm512_set1_ph((_Float16)-4.97f);
  auto xh = _mm512_mul_ph(half, x);
  auto x2 = _mm512_mul_ph(xh, xh);
  auto t1n = _mm512_add_ph(x2, c1);
  auto t2n = _mm512_fmadd_ph(t1n, x2, c0);
  auto nom = _mm512_mul_ph(t2n, xh);
  auto t1d = _mm512_fmadd_ph(c3d, x2, c2d);
  auto t2d = _mm512_fmadd_ph(t1d, x2, c1d);
  auto den = _mm512_fmadd_ph(t2d, x2, c0);
  auto rcp = _mm512_rcp_ph(den);
  auto mh = _mm512_cmp_ph_mask(xh, hi, _CMP_GT_OQ), ml = _mm512_cmp_ph_mask(xh, lo, _CMP_LT_OQ);
  auto r = _mm512_mul_ph(nom, rcp);
  r = _mm512_mask_blend_ph(mh, r, one);
  r = _mm512_mask_blend_ph(ml, r, _mm512_set1_ph((_Float16)-1.0f));
  r = _mm512_add_ph(r, one);
  return _mm512_mul_ph(r, half);
}
static inline __m512h cvt32(const float* p) {  // 32 fp32 -> 32 fp16, RNE
  __m256i a = _mm512_cvtps_ph(_mm512_loadu_ps(p), _MM_FROUND_NO_EXC);
  __m256i b = _mm512_cvtps_ph(_mm512_loadu_ps(p + 16), _MM_FROUND_NO_EXC);
  return _mm512_castsi512_ph(_mm512_inserti64x4(_mm512_castsi256_si512(a), b, 1));
}
static inline void quant_store(int8_t* out, __m512h h, float scale) {
  auto vs = _mm512_set1_ph((_Float16)scale);
  auto m = _mm512_roundscale_ph(_mm512_mul_ph(h, vs), _MM_FROUND_TO_NEAREST_INT);
  auto c1 = _mm512_min_ph(m, _mm512_set1_ph((_Float16)127.f));
  auto c2 = _mm512_max_ph(c1, _mm512_set1_ph((_Float16)-128.f));
  _mm512_mask_cvtepi16_storeu_epi8(out, 0xffffffff, _mm512_cvtph_epi16(c2));
}
static void dump(const char* dir, const char* name, const void* p, size_t bytes) {
  char path[512]; snprintf(path, sizeof path, "%s/%s.bin", dir, name);
  FILE* f = fopen(path, "wb"); fwrite(p, 1, bytes, f); fclose(f);
}
int main(int argc, char** argv) {
  const char* dir = argc > 1 ? argv[1] : ".";
  // 1. full postop on random rows: N rows x H units, several (in_scale,out_scale) and gate spreads
  const int N = 48, H = 64;
  std::mt19937 rng(20261104);
  std::vector<float> gates((size_t)N * 4 * H);
  std::vector<uint16_t> c((size_t)N * H), c_out, ybf((size_t)N * H);
  std::vector<int8_t> hq((size_t)N * H), yq((size_t)N * H);
  for (int n = 0; n < N; ++n) {
    float spread = (n % 4 == 0) ? 0.5f : (n % 4 == 1) ? 2.0f : (n % 4 == 2) ? 6.0f : 14.0f;
    std::normal_distribution<float> nd(0.f, spread);
    for (int j = 0; j < 4 * H; ++j) gates[(size_t)n * 4 * H + j] = nd(rng);
    std::normal_distribution<float> cd(0.f, (n % 3) + 0.3f);
    for (int j = 0; j < H; ++j) { _Float16 v = (_Float16)cd(rng); memcpy(&c[(size_t)n * H + j], &v, 2); }
  }
  const float in_scale = 127.0f / 0.9731f, out_scale = 127.0f / 1.0f;
  dump(dir, "postop_gates_f32", gates.data(), gates.size() * 4);
  dump(dir, "postop_c_in_f16", c.data(), c.size() * 2);
  c_out = c;
  for (int n = 0; n < N; ++n) {
    const float* g = &gates[(size_t)n * 4 * H];
    for (int j = 0; j < H; j += 32) {
      auto it = sigmoid_ph(cvt32(g + j)), ft = sigmoid_ph(cvt32(g + H + j));
      auto gt = tanh_ph(cvt32(g + 2 * H + j)), ot = sigmoid_ph(cvt32(g + 3 * H + j));
      auto cc = _mm512_loadu_ph(&c_out[(size_t)n * H + j]);
      auto a = _mm512_mul_ph(ft, cc);
      auto cn = _mm512_fmadd_ph(it, gt, a);
      _mm512_storeu_ph(&c_out[(size_t)n * H + j], cn);
      auto ht = _mm512_mul_ph(ot, tanh_ph(cn));
      // fp16 -> fp32 -> bf16 (RNE), lstm_postop_tpp.cpp:66-74
      __m512i hbits = _mm512_castph_si512(ht);
      auto lo = _mm512_cvtph_ps(_mm512_castsi512_si256(hbits));
      auto hi = _mm512_cvtph_ps(_mm512_extracti64x4_epi64(hbits, 1));
      auto bf = _mm512_cvtne2ps_pbh(hi, lo);
      _mm512_storeu_si512(&ybf[(size_t)n * H + j], (__m512i)bf);
      quant_store(&yq[(size_t)n * H + j], ht, out_scale);
      quant_store(&hq[(size_t)n * H + j], ht, in_scale);
    }
  }
  float scales[2] = {in_scale, out_scale};
  dump(dir, "postop_scales_f32", scales, sizeof scales);
  dump(dir, "postop_c_out_f16", c_out.data(), c_out.size() * 2);
  dump(dir, "postop_ybf_bf16", ybf.data(), ybf.size() * 2);
  dump(dir, "postop_yq_i8", yq.data(), yq.size());
  dump(dir, "postop_hq_i8", hq.data(), hq.size());
  // 2. tanh_ph over EVERY binary16 input (65536), sigmoid_ph likewise
  std::vector<uint16_t> all(65536), th(65536), sg(65536);
  for (int i = 0; i < 65536; ++i) all[i] = (uint16_t)i;
  for (int i = 0; i < 65536; i += 32) {
    auto x = _mm512_loadu_ph(&all[i]);
    _mm512_storeu_ph(&th[i], tanh_ph(x));
    _mm512_storeu_ph(&sg[i], sigmoid_ph(x));
  }
  dump(dir, "tanh_all_f16", th.data(), th.size() * 2);
  dump(dir, "sigmoid_all_f16", sg.data(), sg.size() * 2);
  // 3. fma on random triples (incl. wide exponent gaps)
  const int M = 1 << 16;
  std::vector<uint16_t> fa(M), fb(M), fc(M), fo(M);
  for (int i = 0; i < M; ++i) { fa[i] = rng() & 0xffff; fb[i] = rng() & 0xffff; fc[i] = rng() & 0xffff; }
  for (int i = 0; i < M; i += 32)
    _mm512_storeu_ph(&fo[i], _mm512_fmadd_ph(_mm512_loadu_ph(&fa[i]), _mm512_loadu_ph(&fb[i]), _mm512_loadu_ph(&fc[i])));
  dump(dir, "fma_a_f16", fa.data(), M * 2); dump(dir, "fma_b_f16", fb.data(), M * 2);
  dump(dir, "fma_c_f16", fc.data(), M * 2); dump(dir, "fma_out_f16", fo.data(), M * 2);
  return 0;
}
