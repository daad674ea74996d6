// gen_amx_bf16_golden.cpp — golden vectors of the bf16 (prediction / joint) tier on REAL AMX-BF16 + AVX-512 hardware.
// Test infrastructure.  The reference's own build cannot be produced offline, but its arithmetic for this tier is three
// hardware operations whose results are not derivable from IEEE rules: TDPBF16PS (the accumulation inside a tile
// product is not a sequence of IEEE fused multiply-adds), VRCP14PS (a table approximation) and the order in which the
// reference chains them.  This program restates that order with the same instructions and runs it on silicon:
//   * tile product: accumulator tile zeroed, then one TDPBF16PS per 32-wide K tile in ascending K
//     (mlperf_plugins/csrc/tpps/tile_product_tpp.hpp:563-581 `_compute_impl_bf16`; B tiles are [16 k-pairs][16 cols][2],
//     the [OC/32, 2, K/32, 16, 32] packing of models/quant_modules.py:210-231);
//   * prediction LSTM cell (lstm_amx.cpp:187-264): G = x.W_ih (i16o32b0), G = (h.W_hh + G) + (b_ih + b_hh)
//     (i16o32b32_accum, tile_product_tpp.hpp:389-411), fp32 gates with VRCP14PS (el_common_intrin.hpp:66-93,154-185),
//     c = fma(i, g, f*c), h = bf16(o * tanh(c)) (tpps/lstm_postop_tpp.cpp:100-163);
//   * joint fc1 (amx_linear.cpp:119-170): y1 = bf16(f.W_trans) (i16o16b0), y = bf16(relu((g.W_pred + float(y1)) + b))
//     (i16o16b32_accum_relu, tile_product_tpp.hpp:427-449);
//   * joint fc2 (amx_linear.cpp:113-117): logits = y.W2 + b2 (i16o32b32, tile_product_tpp.hpp:346-365).
// Inputs come from a counter-based hash that tests/test_amx_golden.py reproduces in numpy, so only the hardware OUTPUTS
// are stored (tests/golden/amx_bf16_hw.npz; recipe tests/golden/make_golden_amx.sh).
#include <immintrin.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct TileCfg { uint8_t palette, start_row; uint8_t rsv[14]; uint16_t colsb[16]; uint8_t rows[16]; };

static inline float bf2f(uint16_t b) { uint32_t u = (uint32_t)b << 16; float f; memcpy(&f, &u, 4); return f; }
static inline uint16_t f2bf(float f) { uint32_t u; memcpy(&u, &f, 4); u += 0x7fffu + ((u >> 16) & 1u); return (uint16_t)(u >> 16); }

// counter-based generator (murmur3 finaliser): value i of stream `seed`, uniform in [-bound, bound)
static inline uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x85ebca6bu; x ^= x >> 13; x *= 0xc2b2ae35u; x ^= x >> 16; return x; }
static inline float uni(uint32_t seed, uint32_t i, float bound) {
  const uint32_t h = mix(i * 2654435761u + seed * 0x9e3779b9u + 12345u);
  const float u = (float)(h >> 8) * (1.0f / 16777216.0f);       // exact: 24-bit integer * 2^-24
  return (2.0f * u - 1.0f) * bound;                               // each step exact or one fp32 rounding, same in numpy
}
static std::vector<uint16_t> gen_bf16(uint32_t seed, size_t n, float bound) {
  std::vector<uint16_t> v(n);
  for (size_t i = 0; i < n; ++i) v[i] = f2bf(uni(seed, (uint32_t)i, bound));
  return v;
}
static std::vector<float> gen_f32(uint32_t seed, size_t n, float bound) {
  std::vector<float> v(n);
  for (size_t i = 0; i < n; ++i) v[i] = uni(seed, (uint32_t)i, bound);
  return v;
}

// C[N][O] (fp32) = A[N][K] . W[O][K]^T with one zeroed accumulator tile per 16x16 block and ascending K tiles.
static void amx_gemm(const uint16_t* A, const uint16_t* W, float* C, int N, int K, int O) {
  alignas(64) uint16_t at[16][32], bt[16][32];
  alignas(64) float ct[16][16];
  for (int n0 = 0; n0 < N; n0 += 16)
    for (int o0 = 0; o0 < O; o0 += 16) {
      _tile_zero(0);
      for (int k0 = 0; k0 < K; k0 += 32) {
        for (int r = 0; r < 16; ++r) memcpy(at[r], A + (size_t)(n0 + r) * K + k0, 64);
        for (int kp = 0; kp < 16; ++kp)
          for (int c = 0; c < 16; ++c) {
            bt[kp][2 * c] = W[(size_t)(o0 + c) * K + k0 + 2 * kp];
            bt[kp][2 * c + 1] = W[(size_t)(o0 + c) * K + k0 + 2 * kp + 1];
          }
        _tile_loadd(1, at, 64);
        _tile_loadd(2, bt, 64);
        _tile_dpbf16ps(0, 1, 2);
      }
      _tile_stored(0, ct, 64);
      for (int r = 0; r < 16; ++r) memcpy(C + (size_t)(n0 + r) * O + o0, ct[r], 64);
    }
}

static inline __m512 pade_ps(__m512 x) {       // shared rational core + clamp (el_common_intrin.hpp:66-93)
  const __m512 c0 = _mm512_set1_ps(4826.25f), c1 = _mm512_set1_ps(618.75f), c2 = _mm512_set1_ps(13.5f);
  const __m512 c3 = _mm512_set1_ps((float)(1.0 / 28.0)), c1d = _mm512_set1_ps(2227.5f), c2d = _mm512_set1_ps(112.5f);
  const __m512 x2 = _mm512_mul_ps(x, x);
  const __m512 t1n = _mm512_fmadd_ps(c3, x2, c2), t2n = _mm512_fmadd_ps(t1n, x2, c1), t3n = _mm512_fmadd_ps(t2n, x2, c0);
  const __m512 nom = _mm512_mul_ps(t3n, x);
  const __m512 t1d = _mm512_add_ps(x2, c2d), t2d = _mm512_fmadd_ps(t1d, x2, c1d), den = _mm512_fmadd_ps(t2d, x2, c0);
  const __m512 rcp = _mm512_rcp14_ps(den);
  const __mmask16 mh = _mm512_cmp_ps_mask(x, _mm512_set1_ps(4.97f), _CMP_GT_OQ);
  const __mmask16 ml = _mm512_cmp_ps_mask(x, _mm512_set1_ps(-4.97f), _CMP_LT_OQ);
  __m512 r = _mm512_mul_ps(nom, rcp);
  r = _mm512_mask_blend_ps(mh, r, _mm512_set1_ps(1.0f));
  return _mm512_mask_blend_ps(ml, r, _mm512_set1_ps(-1.0f));
}
static inline __m512 tanh_ps(__m512 x) { return pade_ps(x); }
static inline __m512 sigmoid_ps(__m512 x) {
  const __m512 half = _mm512_set1_ps(0.5f);
  __m512 r = pade_ps(_mm512_mul_ps(half, x));
  r = _mm512_add_ps(r, _mm512_set1_ps(1.0f));
  return _mm512_mul_ps(r, half);
}

template <class T>
static void dump(const std::string& dir, const char* name, const std::vector<T>& v) {
  FILE* f = fopen((dir + "/" + name + ".bin").c_str(), "wb");
  if (!f || fwrite(v.data(), sizeof(T), v.size(), f) != v.size()) { perror(name); exit(2); }
  fclose(f);
}

int main(int argc, char** argv) {
  const std::string dir = argc > 1 ? argv[1] : ".";
  if (syscall(SYS_arch_prctl, 0x1023 /* ARCH_REQ_XCOMP_PERM */, 18 /* XFEATURE_XTILEDATA */)) { fprintf(stderr, "AMX permission denied\n"); return 1; }
  TileCfg cfg; memset(&cfg, 0, sizeof cfg); cfg.palette = 1;
  for (int i = 0; i < 8; ++i) { cfg.colsb[i] = 64; cfg.rows[i] = 16; }
  _tile_loadconfig(&cfg);

  const int N = 32, P = 320, H = 1024, J = 512, V = 32;
  // ---- prediction LSTM cell (one layer): seeds 1..7
  auto x = gen_bf16(1, (size_t)N * P, 2.0f), h = gen_bf16(2, (size_t)N * P, 1.0f);
  auto c = gen_f32(3, (size_t)N * P, 2.0f);
  auto w_ih = gen_bf16(4, (size_t)4 * P * P, 0.12f), w_hh = gen_bf16(5, (size_t)4 * P * P, 0.12f);
  auto bias = gen_f32(6, (size_t)4 * P, 0.5f);
  std::vector<float> g_ih((size_t)N * 4 * P), g_hh((size_t)N * 4 * P), gates((size_t)N * 4 * P), cy((size_t)N * P);
  std::vector<uint16_t> hy((size_t)N * P);
  amx_gemm(x.data(), w_ih.data(), g_ih.data(), N, P, 4 * P);
  amx_gemm(h.data(), w_hh.data(), g_hh.data(), N, P, 4 * P);
  for (size_t i = 0; i < gates.size(); i += 16) {      // float_out_accum: (i0 + a0) + b0
    const __m512 s = _mm512_add_ps(_mm512_loadu_ps(&g_hh[i]), _mm512_loadu_ps(&g_ih[i]));
    _mm512_storeu_ps(&gates[i], _mm512_add_ps(s, _mm512_loadu_ps(&bias[i % (4 * P)])));
  }
  for (int n = 0; n < N; ++n)
    for (int j = 0; j < P; j += 16) {
      const float* gr = &gates[(size_t)n * 4 * P];
      const __m512 it = sigmoid_ps(_mm512_loadu_ps(gr + j)), ft = sigmoid_ps(_mm512_loadu_ps(gr + P + j));
      const __m512 gt = tanh_ps(_mm512_loadu_ps(gr + 2 * P + j)), ot = sigmoid_ps(_mm512_loadu_ps(gr + 3 * P + j));
      const __m512 cn = _mm512_fmadd_ps(it, gt, _mm512_mul_ps(ft, _mm512_loadu_ps(&c[(size_t)n * P + j])));
      _mm512_storeu_ps(&cy[(size_t)n * P + j], cn);
      const __m512 ht = _mm512_mul_ps(ot, tanh_ps(cn));
      const __m256i hb = (__m256i)_mm512_cvtneps_pbh(ht);
      _mm256_storeu_si256((__m256i*)&hy[(size_t)n * P + j], hb);
    }
  // ---- joint fc1: seeds 11..15
  auto f = gen_bf16(11, (size_t)N * H, 1.0f), g = gen_bf16(12, (size_t)N * P, 1.0f);
  auto w_trans = gen_bf16(13, (size_t)J * H, 0.1f), w_pred = gen_bf16(14, (size_t)J * P, 0.1f);
  auto b1 = gen_f32(15, (size_t)J, 0.5f);
  std::vector<float> a1((size_t)N * J), a2((size_t)N * J);
  std::vector<uint16_t> y1((size_t)N * J), y((size_t)N * J);
  amx_gemm(f.data(), w_trans.data(), a1.data(), N, H, J);
  for (size_t i = 0; i < a1.size(); i += 16)
    _mm256_storeu_si256((__m256i*)&y1[i], (__m256i)_mm512_cvtneps_pbh(_mm512_loadu_ps(&a1[i])));
  amx_gemm(g.data(), w_pred.data(), a2.data(), N, P, J);
  for (size_t i = 0; i < a2.size(); i += 16) {        // bf16_out_accum_relu: max((i0 + a0) + b0, 0)
    const __m512 prev = _mm512_cvtpbh_ps((__m256bh)_mm256_loadu_si256((const __m256i*)&y1[i]));
    __m512 o = _mm512_add_ps(_mm512_add_ps(_mm512_loadu_ps(&a2[i]), prev), _mm512_loadu_ps(&b1[i % J]));
    o = _mm512_max_ps(o, _mm512_setzero_ps());
    _mm256_storeu_si256((__m256i*)&y[i], (__m256i)_mm512_cvtneps_pbh(o));
  }
  // ---- joint fc2: seeds 21..22, input = the hardware y above
  auto w2 = gen_bf16(21, (size_t)V * J, 0.2f);
  auto b2 = gen_f32(22, (size_t)V, 0.5f);
  std::vector<float> logits((size_t)N * V);
  amx_gemm(y.data(), w2.data(), logits.data(), N, J, V);
  for (size_t i = 0; i < logits.size(); i += 16)      // float_out: i0 + b0
    _mm512_storeu_ps(&logits[i], _mm512_add_ps(_mm512_loadu_ps(&logits[i]), _mm512_loadu_ps(&b2[i % V])));
  _tile_release();

  dump(dir, "lstm_g_ih_f32", g_ih); dump(dir, "lstm_gates_f32", gates); dump(dir, "lstm_hy_bf16", hy); dump(dir, "lstm_cy_f32", cy);
  dump(dir, "fc1_acc_trans_f32", a1); dump(dir, "fc1_y1_bf16", y1); dump(dir, "fc1_y_bf16", y); dump(dir, "fc2_logits_f32", logits);
  // two spot values so the numpy twin of the generator can be checked before anything else
  printf("x[7]=%04x w_ih[12345]=%04x bias[100]=%a c[999]=%a\n", x[7], w_ih[12345], bias[100], c[999]);
  return 0;
}
