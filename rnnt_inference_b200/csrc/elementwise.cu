// elementwise.cu — the HBM-bound byte kernels around the encoder: input quantisation + K padding,
// stack_time, and the (CUDA-core, order-defined) bf16 linear used to hoist Joint's transcription half.
#include "kernels.cuh"
#include "gate_math.cuh"

namespace rnnt {

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

// y[m][o] = bf16( sum_k x[m][k] * w[o][k] ), fp32 FMA sequential in k from 0 (oracle order; i16o16b0,
// amx_linear.cpp:137-150).  CTA tile 64 rows x 64 outputs; one thread owns a 4x4 block of outputs (16 independent
// chains), operands staged in shared memory as fp32 in k-chunks of 32.
__global__ void __launch_bounds__(256)
linear_bf16_bf16_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w, uint16_t* __restrict__ y,
                        long long M, int K, int O) {
  __shared__ float xs[LB_TK][LB_TM + 4];
  __shared__ float ws[LB_TK][LB_TO + 4];
  const long long m0 = (long long)blockIdx.y * LB_TM;
  const int o0 = blockIdx.x * LB_TO;
  const int tm = (threadIdx.x >> 4) << 2;   // 0..60
  const int to = (threadIdx.x & 15) << 2;   // 0..60
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += LB_TK) {
    // stage: 64 rows x 32 k of x and of w (each thread 8 elements of each)
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
      if (o < O) y[m * O + o] = (uint16_t)float_to_bf16_bits(acc[i][j]);
    }
  }
}

}  // namespace rnnt
