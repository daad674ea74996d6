// frontend.cu — the audio front-end in front of the RNN-T hot path, for sm_100a.
//
// Reference: datasets/parts/features.py:186-252 (FilterbankFeatures.forward) and the plugin operators it calls
// (mlperf_plugins/csrc/{preemphasis,power_spectrum,frame_splicing,normalization}.cpp).  Operators here:
//   preemphasis_kernel      y = x[i] - 0.97 x[i-1], reflect-padded by n_fft/2 (preemphasis.cpp:10-46)
//   logmel_kernel           torch.stft(512, hop 160, hann 320, center=False) + power_spectrum + dither^2 + mel filterbank +
//                           log (features.py:201-227) in ONE kernel: a warp runs the 512-point real FFT of a frame in
//                           shared memory (256-point complex FFT + split), the spectrum never goes to HBM
//   power_spectrum_kernel   the stand-alone operator (power_spectrum.cpp:10-33), for callers that bring their own STFT
//   frame_splicing_kernel   [N, C, T] -> [N, 3C, ceil(T/3)] (frame_splicing.cpp:12-45)
//   layernorm_pad_kernel    per-(sample, feature) normalisation over time with unbiased variance, output padded in batch
//                           and features (normalization.cpp:10-64, i_layernorm_tpp.cpp:10-76)
// All of it is HBM-bound byte shuffling except the FFT (about 12 K flop per frame on CUDA cores; fp32 like the reference); the
// mel product only visits the non-zero span of each triangular filter (~520 FMA per frame instead of 80 x 257).  Floating point: tolerances are stated in tests/test_frontend.py.
#include <algorithm>

#include "kernels.cuh"

namespace rnnt {

__global__ void preemphasis_kernel(const float* __restrict__ x, const int32_t* __restrict__ lens, float* __restrict__ out,
                                   int N, int L, int pad, float coeff) {
  const int n = blockIdx.y;
  const int lp = L + 2 * pad;
  const int ln = lens[n];
  const float* xr = x + (size_t)n * L;
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < lp; j += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (j < ln + 2 * pad && ln > 0) {
      int i = j - pad;
      if (i < 0) i = -i;                       // reflect, edge not repeated
      if (i >= ln) i = 2 * (ln - 1) - i;
      i = max(0, min(i, ln - 1));              // ln <= pad: the reference is undefined there; stay inside the row
      const float a = xr[i], b = i > 0 ? xr[i - 1] : 0.f;
      v = __fmaf_rn(-coeff, b, a);             // a - coeff * b contracted, as the AVX-512 build of the reference does
    }
    out[(size_t)n * lp + j] = v;
  }
}

__global__ void power_spectrum_kernel(const float2* __restrict__ x, const int32_t* __restrict__ lens, float* __restrict__ out,
                                      int N, int T, int F) {
  const long long total = (long long)N * T * F;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(e / ((long long)T * F));
    const int t = (int)((e / F) % T);
    float v = 0.f;
    if (t < lens[n]) { const float2 c = x[e]; v = __fadd_rn(__fmul_rn(c.x, c.x), __fmul_rn(c.y, c.y)); }
    out[e] = v;
  }
}

__global__ void frame_splicing_kernel(const float* __restrict__ x, const int32_t* __restrict__ lens, float* __restrict__ out,
                                      int N, int C, int T, int factor) {
  const int T2 = (T + factor - 1) / factor;
  const long long total = (long long)N * C * factor * T2;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(e % T2);
    const int c2 = (int)((e / T2) % (C * factor));
    const int n = (int)(e / ((long long)T2 * C * factor));
    const int k = c2 / C, c = c2 % C;
    const int src = j * factor + k;
    out[e] = (src < lens[n] && src < T) ? x[((size_t)n * C + c) * T + src] : 0.f;
  }
}

// one warp per (sample, feature) row: sums in fp32 like the reference's 16-lane accumulators, two passes over the row
__global__ void layernorm_pad_kernel(const float* __restrict__ x, const int32_t* __restrict__ lens, const float* __restrict__ weight,
                                     const float* __restrict__ bias, float* __restrict__ out, int N, int C, int T, int n_pad,
                                     int c_out, int t_out, float eps, int unbiased) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int rows = n_pad * c_out;
  if (warp >= rows) return;
  const int n = warp / c_out, c = warp % c_out;
  float* o = out + (size_t)warp * t_out;
  const int ln = (n < N && c < C) ? min(lens[n], min(T, t_out)) : 0;
  if (ln > 0) {
    const float* r = x + ((size_t)n * C + c) * T;
    float s = 0.f, s2 = 0.f;
    for (int j = lane; j < ln; j += 32) { const float v = r[j]; s += v; s2 = fmaf(v, v, s2); }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, off); s2 += __shfl_xor_sync(0xffffffffu, s2, off); }
    const float mean = s / (float)ln, mean2 = s2 / (float)ln;
    float var = mean2 - mean * mean + eps;
    if (unbiased && ln > 1) var *= (float)ln / (float)(ln - 1);
    const float rstd = 1.0f / sqrtf(var);
    for (int j = lane; j < ln; j += 32) {
      const float w = weight ? weight[(size_t)c * T + j] : 1.f, b = bias ? bias[(size_t)c * T + j] : 0.f;
      o[j] = (r[j] - mean) * (w * rstd) + b;
    }
  }
  for (int j = ln + lane; j < t_out; j += 32) o[j] = 0.f;        // the reference leaves these unwritten
}

// ---------------------------------------------------------------------------------------------- STFT + power + mel + log
constexpr int FE_NFFT = 512, FE_BINS = 257, FE_WARPS = 8, FE_TF = 8;     // frames per warp task (one 32-byte sector per output row)
constexpr int FE_MAX_MELS = 96;
constexpr int FE_MAXW = 64;                                              // widest filter (bins) kept in shared memory

struct LogmelSmem {
  static constexpr int FB_PITCH = FE_MAXW + 1;                           // [mel][bin - first bin of the mel], odd pitch: no conflicts
  static constexpr int FB_BYTES = FE_MAX_MELS * FB_PITCH * 4;            // 24,960
  static constexpr int RANGE_BYTES = FE_MAX_MELS * 8;                    // (first bin, bins) per mel
  static constexpr int CONST_BYTES = FE_MAX_MELS * 4;                    // log-mel of a frame past the utterance (power = 0)
  static constexpr int TW_BYTES = FE_BINS * 8;                           // exp(-2 pi i k / 512), k = 0..256
  static constexpr int WIN_BYTES = FE_NFFT * 4;                          // window zero-padded to n_fft
  static constexpr int SHARED_BYTES = FB_BYTES + RANGE_BYTES + CONST_BYTES + TW_BYTES + WIN_BYTES;
  static constexpr int WARP_BYTES = (256 * 2 + FE_BINS + 3) * 4 + FE_MAX_MELS * FE_TF * 4;   // z re | z im | power | staged output
  static constexpr int TOTAL = SHARED_BYTES + FE_WARPS * WARP_BYTES + 64;                    // 79,584: two blocks per SM
};
static_assert(LogmelSmem::SHARED_BYTES % 8 == 0 && LogmelSmem::FB_BYTES % 8 == 0, "float2 tables are 8-byte aligned");

__device__ __forceinline__ int bitrev8(int v) { return (int)(__brev((unsigned)v) >> 24); }

// mel m of one frame: sum over the filter's bins in ascending order (= the dense product without its exact zeros)
__device__ __forceinline__ float mel_dot(const float* w, const float* pw, int first, int bins) {
  float acc = 0.f;
  for (int i = 0; i < bins; ++i) acc = fmaf(w[i], pw[first + i], acc);
  return acc;
}

__global__ void __launch_bounds__(FE_WARPS * 32, 2)
logmel_kernel(const float* __restrict__ y, const int32_t* __restrict__ frame_lens, const float* __restrict__ window,
              const float* __restrict__ fb, float* __restrict__ out, int N, int Lp, int T, int hop, int win_len, int n_mels,
              float dither2, float log_bias) {
  extern __shared__ __align__(16) uint8_t fe_smem[];
  float* fbs = reinterpret_cast<float*>(fe_smem);
  int2* range = reinterpret_cast<int2*>(fe_smem + LogmelSmem::FB_BYTES);
  float* cst = reinterpret_cast<float*>(fe_smem + LogmelSmem::FB_BYTES + LogmelSmem::RANGE_BYTES);
  float2* tw = reinterpret_cast<float2*>(fe_smem + LogmelSmem::FB_BYTES + LogmelSmem::RANGE_BYTES + LogmelSmem::CONST_BYTES);
  float* win = reinterpret_cast<float*>(fe_smem + LogmelSmem::SHARED_BYTES - LogmelSmem::WIN_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* zr = reinterpret_cast<float*>(fe_smem + LogmelSmem::SHARED_BYTES + warp * LogmelSmem::WARP_BYTES);
  float* zi = zr + 256;
  float* pw = zi + 256;
  float* stage = pw + FE_BINS + 3;

  // ---- per block: the non-zero span of every mel filter (triangles: 2 .. ~20 bins), its weights, twiddles, window
  for (int m = threadIdx.x; m < FE_MAX_MELS; m += blockDim.x) {
    int first = 0, last = -1;
    if (m < n_mels) {
      const float* row = fb + (size_t)m * FE_BINS;
      first = FE_BINS;
      for (int k = 0; k < FE_BINS; ++k)
        if (row[k] != 0.f) { if (first == FE_BINS) first = k; last = k; }
      if (last < 0) first = 0;
    }
    range[m] = make_int2(first, last - first + 1);
  }
  for (int k = threadIdx.x; k < FE_BINS; k += blockDim.x) {
    float s, c;
    sincospif((float)k / 256.0f, &s, &c);
    tw[k] = make_float2(c, -s);
  }
  const int left = (FE_NFFT - win_len) / 2;                   // torch.stft centres a short window inside n_fft
  for (int j = threadIdx.x; j < FE_NFFT; j += blockDim.x) win[j] = (j >= left && j < left + win_len) ? window[j - left] : 0.f;
  __syncthreads();
  for (int e = threadIdx.x; e < FE_MAX_MELS * FE_MAXW; e += blockDim.x) {
    const int m = e / FE_MAXW, i = e % FE_MAXW;
    const int2 r = range[m];
    fbs[m * LogmelSmem::FB_PITCH + i] = (m < n_mels && i < r.y && r.y <= FE_MAXW) ? fb[(size_t)m * FE_BINS + r.x + i] : 0.f;
  }
  __syncthreads();
  // a filter wider than FE_MAXW bins (not a mel triangle at these sizes) is read from global memory instead
  auto weights = [&](int m, int2 r) { return r.y <= FE_MAXW ? fbs + m * LogmelSmem::FB_PITCH : fb + (size_t)m * FE_BINS + r.x; };
  for (int k = lane; k < FE_BINS; k += 32) pw[k] = dither2;    // a frame past the utterance: power_spectrum zeroed it, + dither^2
  __syncwarp();
  if (warp == 0)
    for (int m = lane; m < n_mels; m += 32) { const int2 r = range[m]; cst[m] = logf(mel_dot(weights(m, r), pw, r.x, r.y) + log_bias); }
  __syncthreads();

  const int groups = (T + FE_TF - 1) / FE_TF;
  const long long tasks = (long long)N * groups;
  for (long long task = (long long)blockIdx.x * FE_WARPS + warp; task < tasks; task += (long long)gridDim.x * FE_WARPS) {
    const int n = (int)(task / groups), t0 = (int)(task % groups) * FE_TF;
    const int flen = frame_lens[n];
    const float* yr = y + (size_t)n * Lp;
    for (int ft = 0; ft < FE_TF; ++ft) {
      const int t = t0 + ft;
      if (t >= T) break;
      if (t >= flen) {
        for (int m = lane; m < n_mels; m += 32) stage[m * FE_TF + ft] = cst[m];
        continue;
      }
      // ---- windowed frame as 256 complex points z[m] = s[2m] + i s[2m+1], stored bit-reversed
      const float* fr = yr + (size_t)t * hop;
      for (int m = lane; m < 256; m += 32) {
        const int j = 2 * m;
        float a = 0.f, b = 0.f;
        if (j + 1 >= left && j < left + win_len) {
          const float2 wv = *reinterpret_cast<const float2*>(win + j);
          a = fr[j] * wv.x; b = fr[j + 1] * wv.y;
        }
        const int r = bitrev8(m);
        zr[r] = a; zi[r] = b;
      }
      __syncwarp();
      // ---- 256-point radix-2 FFT (decimation in time), 128 butterflies per stage
#pragma unroll 1
      for (int half = 1; half < 256; half <<= 1) {
        const int tstep = 256 / half;                         // twiddle exp(-2 pi i pos / (2 half)) = tw[pos * 256 / half]
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int b = lane + 32 * q;
          const int pos = b & (half - 1);
          const int i0 = ((b - pos) << 1) + pos, i1 = i0 + half;
          const float2 w = tw[pos * tstep];
          const float xr = zr[i1], xi = zi[i1];
          const float tr = w.x * xr - w.y * xi, ti = w.x * xi + w.y * xr;
          const float ar = zr[i0], ai = zi[i0];
          zr[i1] = ar - tr; zi[i1] = ai - ti;
          zr[i0] = ar + tr; zi[i0] = ai + ti;
        }
        __syncwarp();
      }
      // ---- split into the 257 bins of the real transform, power
      for (int k = lane; k < FE_BINS; k += 32) {
        const int ka = k & 255, kb = (256 - k) & 255;
        const float ar = zr[ka], ai = zi[ka], br = zr[kb], bi = -zi[kb];          // Z[k], conj(Z[256 - k])
        const float er = 0.5f * (ar + br), ei = 0.5f * (ai + bi);                 // even part
        const float dr = 0.5f * (ar - br), di = 0.5f * (ai - bi);                 // (Z - conj Z') / 2; odd part = -i * that
        const float orr = di, oi = -dr;
        const float2 w = tw[k];
        const float xr = er + (w.x * orr - w.y * oi), xi = ei + (w.x * oi + w.y * orr);
        pw[k] = xr * xr + xi * xi + dither2;
      }
      __syncwarp();
      // ---- mel filterbank + log: lane = mel m, m + 32, m + 64
      for (int m = lane; m < n_mels; m += 32) {
        const int2 r = range[m];
        stage[m * FE_TF + ft] = logf(mel_dot(weights(m, r), pw, r.x, r.y) + log_bias);
      }
      __syncwarp();
    }
    __syncwarp();
    // ---- write the staged [mel][FE_TF] block: eight consecutive frames of a mel row are one 32-byte sector
    const int nt = min(FE_TF, T - t0);
    for (int e = lane; e < n_mels * FE_TF; e += 32) {
      const int m = e / FE_TF, ft = e % FE_TF;
      if (ft < nt) out[((size_t)n * n_mels + m) * T + t0 + ft] = stage[e];
    }
    __syncwarp();
  }
}

}  // namespace rnnt

// ---------------------------------------------------------------------------------------------- C ABI
#include "../../include/rnnt_b200.h"
#include "api_internal.h"

using namespace rnnt;

extern "C" {

int rnnt_preemphasis_f32(const float* x, const int32_t* lens, float coeff, int pad, float* out, int N, int L, void* stream) {
  if (!x || !lens || !out || N <= 0 || L <= 0 || pad < 0) return api_fail(RNNT_ERR_INVALID, "preemphasis: bad arguments (N=%d L=%d pad=%d)", N, L, pad);
  const int lp = L + 2 * pad;
  dim3 grid(std::min((lp + 255) / 256, 64), N);
  preemphasis_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, lens, out, N, L, pad, coeff);
  if (cudaGetLastError() != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "preemphasis launch failed");
  return RNNT_OK;
}

int rnnt_power_spectrum_f32(const float* x, const int32_t* lens, float* out, int N, int T, int F, void* stream) {
  if (!x || !lens || !out || N <= 0 || T <= 0 || F <= 0 || (reinterpret_cast<uintptr_t>(x) & 7))
    return api_fail(RNNT_ERR_INVALID, "power_spectrum: bad arguments (N=%d T=%d F=%d; input 8-byte aligned)", N, T, F);
  const long long total = (long long)N * T * F;
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  power_spectrum_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(x), lens, out, N, T, F);
  if (cudaGetLastError() != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "power_spectrum launch failed");
  return RNNT_OK;
}

int rnnt_frame_splicing_f32(const float* x, const int32_t* lens, float* out, int N, int C, int T, int factor, void* stream) {
  if (!x || !lens || !out || N <= 0 || C <= 0 || T <= 0 || factor <= 0)
    return api_fail(RNNT_ERR_INVALID, "frame_splicing: bad arguments (N=%d C=%d T=%d factor=%d)", N, C, T, factor);
  const long long total = (long long)N * C * factor * ((T + factor - 1) / factor);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  frame_splicing_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, lens, out, N, C, T, factor);
  if (cudaGetLastError() != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "frame_splicing launch failed");
  return RNNT_OK;
}

int rnnt_layernorm_pad_f32(const float* x, const int32_t* lens, const float* weight, const float* bias, float* out, int N, int C,
                           int T, int n_pad, int c_out, int t_out, float eps, int unbiased, void* stream) {
  if (!x || !lens || !out || N <= 0 || C <= 0 || T <= 0 || n_pad < N || c_out < C || t_out <= 0)
    return api_fail(RNNT_ERR_INVALID, "layernorm_pad: bad arguments (N=%d C=%d T=%d n_pad=%d c_out=%d t_out=%d)", N, C, T, n_pad, c_out, t_out);
  const long long rows = (long long)n_pad * c_out;
  const int blocks = (int)((rows * 32 + 255) / 256);
  layernorm_pad_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, lens, weight, bias, out, N, C, T, n_pad, c_out, t_out, eps, unbiased);
  if (cudaGetLastError() != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "layernorm_pad launch failed");
  return RNNT_OK;
}

int rnnt_logmel_f32(const float* y, const int32_t* frame_lens, const float* window, const float* fb, float* out, int N, int Lp,
                    int T, int n_fft, int hop, int win_len, int n_mels, float dither2, float log_bias, void* stream) {
  if (!y || !frame_lens || !window || !fb || !out || N <= 0 || T <= 0 || hop <= 0)
    return api_fail(RNNT_ERR_INVALID, "logmel: bad arguments");
  if (n_fft != FE_NFFT || win_len <= 0 || win_len > n_fft || (win_len & 1) || n_mels <= 0 || n_mels > FE_MAX_MELS)
    return api_fail(RNNT_ERR_INVALID, "logmel: n_fft must be 512, win_len even and <= 512, n_mels <= %d (got n_fft=%d win_len=%d n_mels=%d)",
                    FE_MAX_MELS, n_fft, win_len, n_mels);
  if ((long long)(T - 1) * hop + n_fft > Lp)
    return api_fail(RNNT_ERR_INVALID, "logmel: %d frames of hop %d need %lld samples per row, rows have %d", T, hop, (long long)(T - 1) * hop + n_fft, Lp);
  cudaError_t e = cudaFuncSetAttribute(logmel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, LogmelSmem::TOTAL);
  if (e != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "logmel: shared-memory opt-in failed: %s", cudaGetErrorString(e));
  const long long tasks = (long long)N * ((T + FE_TF - 1) / FE_TF);
  const int blocks = (int)std::min<long long>((tasks + FE_WARPS - 1) / FE_WARPS, 2 * 148);
  logmel_kernel<<<blocks, FE_WARPS * 32, LogmelSmem::TOTAL, (cudaStream_t)stream>>>(y, frame_lens, window, fb, out, N, Lp, T, hop,
                                                                                    win_len, n_mels, dither2, log_bias);
  if (cudaGetLastError() != cudaSuccess) return api_fail(RNNT_ERR_CUDA, "logmel launch failed");
  return RNNT_OK;
}

}  // extern "C"
