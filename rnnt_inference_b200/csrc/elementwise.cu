// elementwise.cu — the HBM-bound byte kernels around the encoder: input quantisation + K padding,
// stack_time, and the (CUDA-core, order-defined) bf16 linear used to hoist Joint's transcription half.
#include "kernels.cuh"
#include "gate_math.cuh"
#include "../../include/rnnt_rcp_ph_lut.h"

namespace rnnt {

__constant__ uint16_t c_rcp_lut_ew[1024] = {RNNT_RCP_PH_LUT_VALUES};

// lstm_postop (init.cpp:126-128; lstm_postop.cpp:14-72 -> lstm_postop_tpp::ref, tpps/lstm_postop_tpp.cpp:12-98) as a
// stand-alone operator: fp32 gates (rows `ld` floats apart, so the four chunks of a [N][4H] tensor work in place) ->
// fp16 gate math with the reference's instruction sequence (Pade [5/6] + VRCPPH table) -> c updated IN PLACE,
// y int8 (out_scale) or bf16, h int8 (in_scale).  One thread = two adjacent units of one row.
__global__ void lstm_postop_fp16_kernel(const float* __restrict__ it, const float* __restrict__ ft, const float* __restrict__ gt,
                                        const float* __restrict__ ot, long long ld, uint16_t* __restrict__ ct, float in_scale,
                                        float out_scale, int skip_quant_y, uint16_t* __restrict__ y_bf16,
                                        int8_t* __restrict__ y_q, int8_t* __restrict__ h_q, int N, int H) {
  __shared__ uint16_t lut[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) lut[i] = c_rcp_lut_ew[i];
  __syncthreads();
  const HalfConsts k = make_half_consts();
  const __half2 in_s = __float2half2_rn(in_scale), out_s = __float2half2_rn(out_scale);
  const int pairs = H >> 1;
  const long long total = (long long)N * pairs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / pairs;
    const int j = (int)(i - n * pairs) * 2;
    const long long g = n * ld + j, o = n * H + j;
    const __half2 ig = sigmoid_ph2(__floats2half2_rn(it[g], it[g + 1]), k, lut);
    const __half2 fg = sigmoid_ph2(__floats2half2_rn(ft[g], ft[g + 1]), k, lut);
    const __half2 gg = tanh_ph2(__floats2half2_rn(gt[g], gt[g + 1]), k, lut);
    const __half2 og = sigmoid_ph2(__floats2half2_rn(ot[g], ot[g + 1]), k, lut);
    const __half2 c_old = bits_h2(*reinterpret_cast<const uint32_t*>(ct + o));
    const __half2 c_new = __hfma2(ig, gg, __hmul2_rn(fg, c_old));
    *reinterpret_cast<uint32_t*>(ct + o) = h2_bits(c_new);
    const __half2 h = __hmul2_rn(og, tanh_ph2(c_new, k, lut));
    *reinterpret_cast<uint16_t*>(h_q + o) = (uint16_t)quant_ph2(h, in_s);
    if (skip_quant_y) {
      const float2 hf = __half22float2(h);
      *reinterpret_cast<uint32_t*>(y_bf16 + o) = float_to_bf16_bits(hf.x) | (float_to_bf16_bits(hf.y) << 16);
    } else {
      *reinterpret_cast<uint16_t*>(y_q + o) = (uint16_t)quant_ph2(h, out_s);
    }
  }
}

// Final h of batch tiles that ran exactly ONE step: the recurrent kernel leaves it in slot 1 of its K-chunk-major
// ping-pong buffer (writing the caller's hx inside the kernel would race with slice CTAs still reading h_0 from it);
// this copies it to hx [N][1024].  grid = batch tiles, bit b of `mask` selects tile b.
__global__ void rec_copy_h1_kernel(const int8_t* __restrict__ h_buf, int8_t* __restrict__ hx, int num_btiles, int N,
                                   unsigned mask) {
  const int b = blockIdx.x;
  if (!((mask >> b) & 1u)) return;
  const uint4* src = reinterpret_cast<const uint4*>(h_buf + ((size_t)(num_btiles + b) * REC_KC) * REC_BM * 128);
  for (int i = threadIdx.x; i < REC_KC * REC_BM * 8; i += blockDim.x) {      // 16-byte pieces: [chunk 8][row 128][8]
    const int kc = i / (REC_BM * 8), row = (i / 8) % REC_BM, part = i % 8;
    const int n = b * REC_BM + row;
    if (n < N) reinterpret_cast<uint4*>(hx + (size_t)n * REC_H + kc * 128)[part] = src[(i & ~7) | (part ^ (row & 7))];   // chunks are stored 128-byte-swizzled
  }
}

// quant_tpp::quant_ps_epi8 (tpps/quant_tpp.cpp:6-17) + zero pad of K (lstm_amx.cpp:48-50).
// One thread produces 4 output bytes; out rows are k_out bytes (k_out % 4 == 0).
__global__ void quantize_pad_kernel(const float* __restrict__ in, int8_t* __restrict__ out, float scale, long long rows,
                                    int k_in, int k_out) {
  const int words_per_row = k_out >> 2;
  const long long total = rows * words_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / words_per_row;
    const int k0 = (int)(i - r * words_per_row) << 2;
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + j;
      int q = 0;
      if (k < k_in) {
        float m = rintf(__fmul_rn(in[r * k_in + k], scale));   // VRNDSCALEPS nearest-even
        m = fminf(m, 127.f);
        m = fmaxf(m, -128.f);
        q = (int)m;
      }
      packed |= ((uint32_t)q & 0xffu) << (8 * j);
    }
    reinterpret_cast<uint32_t*>(out)[i] = packed;
  }
}

// Ragged variant for the SUT: the batch arrives as the unpadded concatenation of its utterances (packed [sum T_i][k_in]
// fp32, utterance n starting at row offsets[n]); this writes the time-major padded int8 tensor {T, N, k_out} that
// pad_sequence + lstm_amx_int8's quantiser would have produced (models/pytorch_sut.py:85-104, lstm_amx.cpp:35-50):
// out[t][n][k] = t < lens[n] && k < k_in ? q(packed[offsets[n] + t][k]) : 0.  Rows n >= n_valid are all zero (batch pad).
__global__ void quantize_ragged_kernel(const float* __restrict__ packed, const long long* __restrict__ offsets,
                                       const int32_t* __restrict__ lens, int8_t* __restrict__ out, float scale, int T, int N,
                                       int k_in, int k_out) {
  const int words_per_row = k_out >> 2;
  const long long total = (long long)T * N * words_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int wi = (int)(i % words_per_row);
    const long long r = i / words_per_row;
    const int n = (int)(r % N), t = (int)(r / N);
    const int k0 = wi << 2;
    uint32_t packed_q = 0;
    if (t < lens[n] && k0 < k_in) {
      const float* src = packed + (offsets[n] + t) * k_in + k0;
      if (k0 + 3 < k_in) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(src));
        const float xs[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float m = rintf(__fmul_rn(xs[j], scale));
          m = fmaxf(fminf(m, 127.f), -128.f);
          packed_q |= ((uint32_t)(int)m & 0xffu) << (8 * j);
        }
      } else {
        for (int j = 0; j < 4 && k0 + j < k_in; ++j) {
          float m = rintf(__fmul_rn(src[j], scale));
          m = fmaxf(fminf(m, 127.f), -128.f);
          packed_q |= ((uint32_t)(int)m & 0xffu) << (8 * j);
        }
      }
    }
    reinterpret_cast<uint32_t*>(out)[i] = packed_q;
  }
}

// stack_time (stack_time.cpp:27-59), 16 bytes per thread.  out[t2][n][(t%f)*C + c] = in[t][n][c] if t < lens[n] else 0.
__global__ void stack_time_kernel(const int8_t* __restrict__ in, const int32_t* __restrict__ lens, int factor,
                                  int8_t* __restrict__ out, int T, int N, int C) {
  const int vec_per_row = C >> 4;
  const int T2 = (T + factor - 1) / factor;
  const long long total = (long long)T2 * factor * N * vec_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vec_per_row);
    long long r = i / vec_per_row;
    const int n = (int)(r % N);
    const int t = (int)(r / N);            // t in [0, T2*factor)
    int4 val = make_int4(0, 0, 0, 0);
    if (t < T && t < lens[n]) val = __ldg(reinterpret_cast<const int4*>(in + ((long long)t * N + n) * C) + v);
    int8_t* dst = out + (((long long)(t / factor) * N + n) * factor + (t % factor)) * C;
    reinterpret_cast<int4*>(dst)[v] = val;
  }
}

// Order-defined bf16 linear: acc[m][o] = sum_k x[m][k] * w[o][k], fp32 FMA sequential in k from 0 (the order the oracle
// defines for the AMX bf16 kernels, tile_product_tpp.hpp:563-581).  CTA tile 64 rows x 64 outputs; one thread owns a
// 4x4 block of outputs (16 independent chains), operands staged in shared memory as fp32 in k-chunks of 32.
// Epilogues (tile_product_tpp.hpp:347-449):
//   LIN_BF16        y_bf16 = bf16(acc)                                        i16o16b0       (amx_linear.cpp:137-150)
//   LIN_F32_BIAS    y_f32  = acc + bias                                       i16o32b32      (amx_linear.cpp:113-117)
//   LIN_F32         y_f32  = acc                                              i16o32b0       (lstm_amx.cpp:218-232)
//   LIN_ACCUM_RELU  y_bf16 = bf16(max((acc + float(prev_bf16)) + bias, 0))    i16o16b32_accum_relu (amx_linear.cpp:152-168)
template <int MODE>
__global__ void __launch_bounds__(256)
linear_bf16_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w, const float* __restrict__ bias,
                   const uint16_t* __restrict__ prev, void* __restrict__ yout, long long M, int K, int O, const RowSkip skip) {
  __shared__ float xs[LB_TK][LB_TM + 4];
  __shared__ float ws[LB_TK][LB_TO + 4];
  const long long m0 = (long long)blockIdx.y * LB_TM;
  const int o0 = blockIdx.x * LB_TO;
  if (skip.dead(m0, LB_TM)) return;
  const int tm = (threadIdx.x >> 4) << 2;   // 0..60
  const int to = (threadIdx.x & 15) << 2;   // 0..60
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += LB_TK) {
    for (int e = threadIdx.x; e < LB_TM * LB_TK; e += 256) {
      const int r = e / LB_TK, kk = e % LB_TK;
      const long long m = m0 + r;
      const int k = k0 + kk;
      xs[kk][r] = (m < M && k < K) ? bf16_bits_to_float(x[m * K + k]) : 0.f;
      const int o = o0 + r;
      ws[kk][r] = (o < O && k < K) ? bf16_bits_to_float(w[(long long)o * K + k]) : 0.f;
    }
    __syncthreads();
    const int kmax = min(LB_TK, K - k0);
    for (int kk = 0; kk < kmax; ++kk) {
      const float4 xv = *reinterpret_cast<const float4*>(&xs[kk][tm]);
      const float4 wv = *reinterpret_cast<const float4*>(&ws[kk][to]);
      const float xa[4] = {xv.x, xv.y, xv.z, xv.w};
      const float wa[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = __fmaf_rn(xa[i], wa[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + tm + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = o0 + to + j;
      if (o >= O) continue;
      const long long idx = m * O + o;
      if constexpr (MODE == LIN_BF16) {
        reinterpret_cast<uint16_t*>(yout)[idx] = (uint16_t)float_to_bf16_bits(acc[i][j]);
      } else if constexpr (MODE == LIN_F32_BIAS) {
        reinterpret_cast<float*>(yout)[idx] = __fadd_rn(acc[i][j], bias[o]);
      } else if constexpr (MODE == LIN_F32) {
        reinterpret_cast<float*>(yout)[idx] = acc[i][j];
      } else {
        const float v = __fadd_rn(__fadd_rn(acc[i][j], bf16_bits_to_float(prev[idx])), bias[o]);
        reinterpret_cast<uint16_t*>(yout)[idx] = (uint16_t)float_to_bf16_bits(fmaxf(v, 0.f));
      }
    }
  }
}
template __global__ void linear_bf16_kernel<LIN_BF16>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_kernel<LIN_F32_BIAS>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_kernel<LIN_F32>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_kernel<LIN_ACCUM_RELU>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);

// ---------------------------------------------------------------------------------------------------------------
// Large-M variant of the same contract (the hoisted f.W_trans product: M = T*N rows, K = 1024, O = 512 is 14 % of a
// bench step).  128 x 128 output tile per CTA, 8 x 8 outputs per thread as four 4 x 4 quadrants (rows tm, tm + 64;
// columns to, to + 64) so every shared-memory read is a conflict-free 16-byte vector and 64 FFMAs follow 4 loads; the
// next K tile travels global -> registers while the current one is multiplied.  Operands are widened to fp32 once,
// when they are staged, and stored k-major with the row index XOR-ed by the k octet so the transposing stores of a
// warp hit 32 different banks.  The fp32 chain of every output is still k = 0, 1, 2, ... in order (bit-exact).
// Requires K % 32 == 0 and 16-byte aligned x, w (api.cu falls back to the 64 x 64 kernel otherwise).
template <int MODE>
__device__ __forceinline__ void lin_store(float acc, long long idx, int o, const float* bias, const uint16_t* prev, void* yout) {
  if constexpr (MODE == LIN_BF16) {
    reinterpret_cast<uint16_t*>(yout)[idx] = (uint16_t)float_to_bf16_bits(acc);
  } else if constexpr (MODE == LIN_F32_BIAS) {
    reinterpret_cast<float*>(yout)[idx] = __fadd_rn(acc, bias[o]);
  } else if constexpr (MODE == LIN_F32) {
    reinterpret_cast<float*>(yout)[idx] = acc;
  } else {
    const float v = __fadd_rn(__fadd_rn(acc, bf16_bits_to_float(prev[idx])), bias[o]);
    reinterpret_cast<uint16_t*>(yout)[idx] = (uint16_t)float_to_bf16_bits(fmaxf(v, 0.f));
  }
}

template <int MODE>
__global__ void __launch_bounds__(256, 2)
linear_bf16_big_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w, const float* __restrict__ bias,
                       const uint16_t* __restrict__ prev, void* __restrict__ yout, long long M, int K, int O, const RowSkip skip) {
  __shared__ __align__(16) float xs[LBB_TK][LBB_T];
  __shared__ __align__(16) float ws[LBB_TK][LBB_T];
  const long long m0 = (long long)blockIdx.y * LBB_T;
  const int o0 = blockIdx.x * LBB_T;
  if (skip.dead(m0, LBB_T)) return;
  const int tm = (threadIdx.x >> 4) << 2;   // 0..60 (and + 64)
  const int to = (threadIdx.x & 15) << 2;   // 0..60 (and + 64)
  const int lrow = threadIdx.x >> 2;        // loader: rows lrow, lrow + 64; k octet lkq
  const int lkq = threadIdx.x & 3;
  uint4 xr[2], wr[2];
  auto gload = [&](int k0) {
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const long long m = m0 + lrow + 64 * p;
      const int o = o0 + lrow + 64 * p;
      xr[p] = (m < M) ? __ldg(reinterpret_cast<const uint4*>(x + m * K + k0 + lkq * 8)) : make_uint4(0, 0, 0, 0);
      wr[p] = (o < O) ? __ldg(reinterpret_cast<const uint4*>(w + (long long)o * K + k0 + lkq * 8)) : make_uint4(0, 0, 0, 0);
    }
  };
  auto sstore = [&]() {
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const int col = (lrow + 64 * p) ^ (lkq * 8);
      const uint32_t xw[4] = {xr[p].x, xr[p].y, xr[p].z, xr[p].w};
      const uint32_t ww[4] = {wr[p].x, wr[p].y, wr[p].z, wr[p].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        xs[lkq * 8 + 2 * j][col] = __uint_as_float(xw[j] << 16);
        xs[lkq * 8 + 2 * j + 1][col] = __uint_as_float(xw[j] & 0xffff0000u);
        ws[lkq * 8 + 2 * j][col] = __uint_as_float(ww[j] << 16);
        ws[lkq * 8 + 2 * j + 1][col] = __uint_as_float(ww[j] & 0xffff0000u);
      }
    }
  };
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  gload(0);
  for (int k0 = 0; k0 < K; k0 += LBB_TK) {
    sstore();
    __syncthreads();
    if (k0 + LBB_TK < K) gload(k0 + LBB_TK);
#pragma unroll 1
    for (int kg = 0; kg < LBB_TK / 8; ++kg) {
      const int f = kg * 8;
#pragma unroll
      for (int k8 = 0; k8 < 8; ++k8) {
        const int kk = f + k8;
        const float4 x0 = *reinterpret_cast<const float4*>(&xs[kk][tm ^ f]);
        const float4 x1 = *reinterpret_cast<const float4*>(&xs[kk][(tm + 64) ^ f]);
        const float4 w0 = *reinterpret_cast<const float4*>(&ws[kk][to ^ f]);
        const float4 w1 = *reinterpret_cast<const float4*>(&ws[kk][(to + 64) ^ f]);
        const float xa[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        const float wa[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = __fmaf_rn(xa[i], wa[j], acc[i][j]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long m = m0 + tm + (i & 3) + ((i >> 2) << 6);
    if (m >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int o = o0 + to + (jh << 6);
      const long long idx = m * O + o;
      if (MODE == LIN_BF16 && o + 3 < O && (O & 3) == 0) {
        uint2 v;
        v.x = float_to_bf16_bits(acc[i][jh * 4 + 0]) | (float_to_bf16_bits(acc[i][jh * 4 + 1]) << 16);
        v.y = float_to_bf16_bits(acc[i][jh * 4 + 2]) | (float_to_bf16_bits(acc[i][jh * 4 + 3]) << 16);
        *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(yout) + idx) = v;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (o + j < O) lin_store<MODE>(acc[i][jh * 4 + j], idx + j, o + j, bias, prev, yout);
      }
    }
  }
}
template __global__ void linear_bf16_big_kernel<LIN_BF16>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_big_kernel<LIN_F32_BIAS>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_big_kernel<LIN_F32>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);
template __global__ void linear_bf16_big_kernel<LIN_ACCUM_RELU>(const uint16_t*, const uint16_t*, const float*, const uint16_t*, void*, long long, int, int, const RowSkip);

// dequant of an int32 accumulator (i8o32b32, tile_product_tpp.hpp:347-367 for int8): (float(acc) + bias[o]) * scale,
// two separately rounded fp32 operations.
__global__ void dequant_i32_f32_kernel(const int32_t* __restrict__ acc, const float* __restrict__ bias, float scale,
                                       float* __restrict__ out, long long total, int O) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float b = bias ? bias[i % O] : 0.f;
    out[i] = __fmul_rn(__fadd_rn(__int2float_rn(acc[i]), b), scale);
  }
}

// lstm_postop_tpp::ref_bf16 (tpps/lstm_postop_tpp.cpp:100-163) preceded by the float_out_accum epilogue
// (tile_product_tpp.hpp:387-408): gate = (gh + gi) + bias.  gi, gh [N][4P] fp32; c_in [N][P]; writes hy bf16, cy fp32.
__global__ void lstm_postop_f32_kernel(const float* __restrict__ gi, const float* __restrict__ gh, const float* __restrict__ bias,
                                       const float* __restrict__ c_in, uint16_t* __restrict__ hy, float* __restrict__ cy,
                                       int N, int P) {
  const long long total = (long long)N * P;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / P), j = (int)(i % P);
    float gate[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const long long gidx = (long long)n * 4 * P + q * P + j;
      gate[q] = __fadd_rn(__fadd_rn(gh[gidx], gi[gidx]), bias[q * P + j]);
    }
    const float it = sigmoid_ps(gate[0]), ft = sigmoid_ps(gate[1]), gt = tanh_ps(gate[2]), ot = sigmoid_ps(gate[3]);
    const float c_new = __fmaf_rn(it, gt, __fmul_rn(ft, c_in[i]));
    cy[i] = c_new;
    hy[i] = (uint16_t)float_to_bf16_bits(__fmul_rn(ot, tanh_ps(c_new)));
  }
}

// greedy_decode_update (greedy_decode_update.cpp:13-96 + tpps/greedy_decode_update_tpp.cpp:9-51): one warp per lane.
// unfinished: device int, incremented once per lane that has not finished (caller zeroes it first).
__global__ void greedy_update_kernel(const long long* __restrict__ symbols, int32_t* symbols_added, int32_t* res, int32_t* res_idx,
                                     const uint16_t* __restrict__ f, const int32_t* __restrict__ f_lens, int32_t* time_idx,
                                     uint16_t* fi, int32_t* pre_g, uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0,
                                     float* pre_cg1, const uint16_t* __restrict__ hg0, const uint16_t* __restrict__ hg1,
                                     const float* __restrict__ cg0, const float* __restrict__ cg1, int N, int L, int C, int P,
                                     int* unfinished) {
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  const int sym = (int)symbols[n];
  const int flen = f_lens[n];
  const bool upd_g = (sym != 28) && (symbols_added[n] != 30) && (flen != 0);
  int ti = time_idx[n];
  __syncwarp();
  if (upd_g) {
    if (lane == 0) {
      const int ri = res_idx[n] + 1;
      res_idx[n] = ri;
      if (ri >= 0 && ri < L) res[(long long)n * L + ri] = sym;
      symbols_added[n] += 1;
      pre_g[n] = sym;
    }
    for (int k = lane; k < P; k += 32) {
      pre_hg0[(long long)n * P + k] = hg0[(long long)n * P + k];
      pre_hg1[(long long)n * P + k] = hg1[(long long)n * P + k];
      pre_cg0[(long long)n * P + k] = cg0[(long long)n * P + k];
      pre_cg1[(long long)n * P + k] = cg1[(long long)n * P + k];
    }
  } else {
    ti += 1;
    if (lane == 0) { symbols_added[n] = 0; time_idx[n] = ti; }
  }
  if (ti < flen) {
    const uint16_t* src = f + ((long long)ti * N + n) * C;
    for (int k = lane; k < C; k += 32) fi[(long long)n * C + k] = src[k];
    if (lane == 0) atomicAdd(unfinished, 1);
  }
}

}  // namespace rnnt
