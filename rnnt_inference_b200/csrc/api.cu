// api.cu — the C ABI of librnnt_b200.so (include/rnnt_b200.h): argument validation, TMA tensor-map
// construction, workspace carving and kernel launches.  No torch types, no allocation of caller-visible
// buffers, no exceptions; errors are codes + a thread-local message.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>

#include "../../include/rnnt_b200.h"
#ifdef RNNT_DEBUG_HOOKS
#include "../../include/rnnt_b200_debug.h"
#endif
#include "kernels.cuh"
#include "api_internal.h"

using namespace rnnt;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}
#define CUDA_TRY(expr)                                                                                  \
  do {                                                                                                  \
    cudaError_t e_ = (expr);                                                                            \
    if (e_ != cudaSuccess) return fail(RNNT_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// rank <= 3 tiled tensor map with 128-byte swizzle; dims/box innermost first; strides in bytes for dims 1..rank-1
int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
              const uint64_t* strides, const uint32_t* box, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(RNNT_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], es[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides[i];
  CUresult r = enc(m, dt, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RNNT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return RNNT_OK;
}

// one int of device-side status per device, owned by the library (4 bytes; never handed to the caller)
int* device_status() {
  static int* ptr[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    if (cudaMalloc(&ptr[dev], 256) != cudaSuccess) return nullptr;
    cudaMemset(ptr[dev], 0, 256);
  }
  return ptr[dev];
}

// Cooperative kernels of this library (recurrent, greedy decode) wait on flags written by their own CTAs, which is safe
// only while all of a grid's CTAs become resident.  Two such grids from DIFFERENT streams, handed to an idle GPU at the same
// time, can each be given part of the SMs and then wait for ever (observed on an 8-GPU box with two compute streams per
// rank: watchdog code 102 on 3 of 8 ranks).  They are therefore ordered per device, whatever stream they come from: each
// waits for the completion of the one launched before it.  Kernels without inter-CTA waits (GEMMs, byte kernels) are not
// ordered and still overlap with them.  On one stream the wait is already implied and costs nothing on the device.
struct CoopOrder {
  std::unique_lock<std::mutex> lock;
  cudaEvent_t ev = nullptr;
  cudaStream_t st = nullptr;
  int begin(cudaStream_t stream) {
    static std::mutex mu;
    static cudaEvent_t evs[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
    lock = std::unique_lock<std::mutex>(mu);
    if (!evs[dev] && cudaEventCreateWithFlags(&evs[dev], cudaEventDisableTiming) != cudaSuccess)
      return fail(RNNT_ERR_CUDA, "cudaEventCreate failed: %s", cudaGetErrorString(cudaGetLastError()));
    ev = evs[dev]; st = stream;
    cudaError_t e = cudaStreamWaitEvent(st, ev, 0);      // an event never recorded is complete
    if (e != cudaSuccess) return fail(RNNT_ERR_CUDA, "cudaStreamWaitEvent failed: %s", cudaGetErrorString(e));
    return RNNT_OK;
  }
  int end() {
    cudaError_t e = cudaEventRecord(ev, st);
    lock.unlock();
    if (e != cudaSuccess) return fail(RNNT_ERR_CUDA, "cudaEventRecord failed: %s", cudaGetErrorString(e));
    return RNNT_OK;
  }
};

int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}
inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }
inline int64_t round_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

}  // namespace

namespace rnnt {
int api_fail(int code, const char* fmt, ...) {      // api_internal.h: for frontend.cu
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}
}  // namespace rnnt

// RowSkip for a ragged batch: rows_per_step rows per timestep, tile_steps_host[b] live steps of 128-row tile b (NULL = all live)
static int make_skip(RowSkip* skip, int64_t M, int rows_per_step, const int32_t* tile_steps_host, int block_rows) {
  memset(skip, 0, sizeof *skip);
  if (!tile_steps_host) return RNNT_OK;
  (void)block_rows;       // any rows_per_step: RowSkip::dead walks a block of rows that straddles tiles or timesteps
  if (rows_per_step <= 0 || M % rows_per_step != 0 || rows_per_step > 32 * 128)
    return fail(RNNT_ERR_INVALID, "ragged row skipping needs 0 < rows_per_step <= 4096 and M %% rows_per_step == 0 (M=%lld, rows_per_step=%d)",
                (long long)M, rows_per_step);
  skip->rows_per_t = rows_per_step;
  for (int b = 0; b < (rows_per_step + 127) / 128; ++b) skip->steps[b] = tile_steps_host[b];
  return RNNT_OK;
}

template <int MODE>
static int linear_launch(const uint16_t* x, const uint16_t* w, const float* bias, const uint16_t* prev, void* y, int64_t M, int K,
                         int O, cudaStream_t st, const RowSkip* skip_in = nullptr) {
  RowSkip skip;
  if (skip_in) skip = *skip_in; else memset(&skip, 0, sizeof skip);
  // enough 128 x 128 tiles to fill the machine twice over -> the large-tile kernel (FMA-bound instead of LDS-bound)
  const long long big_tiles = ((M + LBB_T - 1) / LBB_T) * ((O + LBB_T - 1) / LBB_T);
  const bool aligned = (((uintptr_t)x | (uintptr_t)w) & 15) == 0 && K % LBB_TK == 0;
  if (aligned && big_tiles >= 2 * 148 && (skip.rows_per_t % LBB_T) == 0) {
    dim3 grid((O + LBB_T - 1) / LBB_T, (unsigned)((M + LBB_T - 1) / LBB_T));
    linear_bf16_big_kernel<MODE><<<grid, 256, 0, st>>>(x, w, bias, prev, y, M, K, O, skip);
  } else {
    dim3 grid((O + LB_TO - 1) / LB_TO, (unsigned)((M + LB_TM - 1) / LB_TM));
    linear_bf16_kernel<MODE><<<grid, 256, 0, st>>>(x, w, bias, prev, y, M, K, O, skip);
  }
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

// ================================================================================================
extern "C" {

const char* rnnt_last_error(void) { return g_err; }
int rnnt_abi_version(void) { return 1; }
int rnnt_device_sm_count(void) { return sm_count(); }

// Synchronises `stream`, then reads and clears the device-side status word of kernels launched by this
// library on the current device.  Returns RNNT_ERR_DEVICE if a device watchdog fired.
int rnnt_check_device_status(void* stream) {
  int* st = device_status();
  if (!st) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  cudaStream_t s = (cudaStream_t)stream;
  int h[8] = {0};       // [0] code, then what the wait that fired left behind: site, CTA, epoch / step
  CUDA_TRY(cudaMemcpyAsync(h, st, sizeof(h), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  if (h[0] != 0) {
    cudaMemsetAsync(st, 0, sizeof(h), s);
    return fail(RNNT_ERR_DEVICE, "device-side watchdog/protocol error %d (101 mbarrier timeout, 102 flag timeout); site %d "
                "(1 decode group barrier, 2 decode rendezvous, 3 recurrent h flags; 0 = an mbarrier) CTA %d epoch/step %d",
                h[0], h[1], h[2], h[3]);
  }
  return RNNT_OK;
}

// Non-synchronising variant for pipelined callers: enqueues a copy of the status word into *host_dst (pinned host memory)
// on `stream`; the caller looks at it after its own synchronisation point and calls rnnt_clear_device_status on error.
int rnnt_peek_device_status(int32_t* host_dst, void* stream) {
  int* st = device_status();
  if (!st || !host_dst) return fail(RNNT_ERR_CUDA, "peek_device_status: no usable CUDA device or null destination");
  CUDA_TRY(cudaMemcpyAsync(host_dst, st, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return RNNT_OK;
}
int rnnt_clear_device_status(void* stream) {
  int* st = device_status();
  if (!st) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  CUDA_TRY(cudaMemsetAsync(st, 0, 8 * sizeof(int), (cudaStream_t)stream));
  return RNNT_OK;
}

int rnnt_quantize_pad_f32_i8(const float* in, int8_t* out, float scale, int64_t rows, int k_in, int k_out, void* stream) {
  if (!in || !out || rows < 0 || k_in <= 0 || k_out < k_in || (k_out & 3) || !aligned(out, 4))
    return fail(RNNT_ERR_INVALID, "quantize_pad: need k_out >= k_in > 0, k_out %% 4 == 0 (got rows=%lld k_in=%d k_out=%d)",
                (long long)rows, k_in, k_out);
  if (rows == 0) return RNNT_OK;
  const long long total = rows * (k_out / 4);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  quantize_pad_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, out, scale, rows, k_in, k_out);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_quantize_ragged_f32_i8(const float* packed, const int64_t* offsets, const int32_t* lens, int8_t* out, float scale, int T,
                                int N, int k_in, int k_out, void* stream) {
  if (!packed || !offsets || !lens || !out || T < 0 || N <= 0 || k_in <= 0 || k_out < k_in || (k_out & 3) || (k_in & 3) ||
      !aligned(out, 4) || !aligned(packed, 16))
    return fail(RNNT_ERR_INVALID, "quantize_ragged: need k_out >= k_in > 0, both %% 4 == 0, 16-byte aligned input (T=%d N=%d k_in=%d k_out=%d)",
                T, N, k_in, k_out);
  if (T == 0) return RNNT_OK;
  const long long total = (long long)T * N * (k_out / 4);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  quantize_ragged_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(packed, reinterpret_cast<const long long*>(offsets), lens, out,
                                                                   scale, T, N, k_in, k_out);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_stack_time_i8(const int8_t* in, const int32_t* lens, int factor, int8_t* out, int T, int N, int C, void* stream) {
  if (!in || !lens || !out || factor <= 0 || T < 0 || N <= 0 || C <= 0 || (C & 15) || !aligned(in, 16) || !aligned(out, 16))
    return fail(RNNT_ERR_INVALID, "stack_time: need C %% 16 == 0 and 16-byte aligned buffers (T=%d N=%d C=%d factor=%d)", T, N,
                C, factor);
  if (T == 0) return RNNT_OK;
  const int T2 = (T + factor - 1) / factor;
  const long long total = (long long)T2 * factor * N * (C / 16);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  stack_time_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, lens, factor, out, T, N, C);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_linear_bf16_bf16_ragged(const uint16_t* x, const uint16_t* w, uint16_t* y, int64_t M, int K, int O, int rows_per_step,
                                 const int32_t* tile_steps_host, void* stream) {
  if (!x || !w || !y || M < 0 || K <= 0 || O <= 0) return fail(RNNT_ERR_INVALID, "linear_bf16_bf16: null pointer or bad size");
  if (M == 0) return RNNT_OK;
  RowSkip skip;
  if (int r = make_skip(&skip, M, rows_per_step, tile_steps_host, LB_TM)) return r;
  return linear_launch<LIN_BF16>(x, w, nullptr, nullptr, y, M, K, O, (cudaStream_t)stream, &skip);
}
int rnnt_linear_bf16_bf16(const uint16_t* x, const uint16_t* w, uint16_t* y, int64_t M, int K, int O, void* stream) {
  if (!x || !w || !y || M < 0 || K <= 0 || O <= 0) return fail(RNNT_ERR_INVALID, "linear_bf16_bf16: bad arguments");
  if (M == 0) return RNNT_OK;
  return rnnt_linear_bf16_bf16_ragged(x, w, y, M, K, O, 0, nullptr, stream);
}

int rnnt_linear_bf16_f32(const uint16_t* x, const uint16_t* w, const float* bias, float* y, int64_t M, int K, int O,
                         void* stream) {
  if (!x || !w || !y || M < 0 || K <= 0 || O <= 0) return fail(RNNT_ERR_INVALID, "linear_bf16_f32: bad arguments");
  if (M == 0) return RNNT_OK;
  if (bias) return linear_launch<LIN_F32_BIAS>(x, w, bias, nullptr, y, M, K, O, (cudaStream_t)stream);
  return linear_launch<LIN_F32>(x, w, nullptr, nullptr, y, M, K, O, (cudaStream_t)stream);
}

int rnnt_linear_bf16_accum_relu(const uint16_t* y1, const uint16_t* x1, const uint16_t* w1, const float* bias, uint16_t* y,
                                int64_t M, int K1, int O, void* stream) {
  if (!y1 || !x1 || !w1 || !bias || !y || M < 0 || K1 <= 0 || O <= 0)
    return fail(RNNT_ERR_INVALID, "linear_bf16_accum_relu: bad arguments");
  if (M == 0) return RNNT_OK;
  return linear_launch<LIN_ACCUM_RELU>(x1, w1, bias, y1, y, M, K1, O, (cudaStream_t)stream);
}

int rnnt_dequant_i32_f32(const int32_t* acc, const float* bias, float scale, float* out, int64_t M, int O, void* stream) {
  if (!acc || !out || M < 0 || O <= 0) return fail(RNNT_ERR_INVALID, "dequant_i32_f32: bad arguments");
  if (M == 0) return RNNT_OK;
  const long long total = M * O;
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  dequant_i32_f32_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(acc, bias, scale, out, total, O);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_lstm_postop_fp16(const float* it, const float* ft, const float* gt, const float* ot, int64_t ld, uint16_t* ct,
                          float in_scale, float out_scale, int skip_quant_y, uint16_t* y_bf16, int8_t* y_q, int8_t* h_q, int N,
                          int H, void* stream) {
  if (!it || !ft || !gt || !ot || !ct || !h_q || (skip_quant_y ? !y_bf16 : !y_q) || N < 0 || H <= 0 || (H & 1) || ld < H || (ld & 1))
    return fail(RNNT_ERR_INVALID, "lstm_postop: null pointer or bad size (N=%d H=%d ld=%lld; H and ld must be even)", N, H, (long long)ld);
  if (!aligned(ct, 4) || !aligned(h_q, 2) || (y_q && !aligned(y_q, 2)) || (y_bf16 && !aligned(y_bf16, 4)))
    return fail(RNNT_ERR_INVALID, "lstm_postop: misaligned buffer");
  if (N == 0) return RNNT_OK;
  const long long total = (long long)N * (H / 2);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 8);
  lstm_postop_fp16_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(it, ft, gt, ot, ld, ct, in_scale, out_scale, skip_quant_y ? 1 : 0,
                                                                    y_bf16, y_q, h_q, N, H);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_lstm_cell_bf16(const uint16_t* x, const uint16_t* h, const float* c, const uint16_t* w_ih, const uint16_t* w_hh,
                        const float* bias, uint16_t* hy, float* cy, int N, int K, int P, float* gates_ws, void* stream) {
  if (!x || !h || !c || !w_ih || !w_hh || !bias || !hy || !cy || !gates_ws || N <= 0 || K <= 0 || P <= 0)
    return fail(RNNT_ERR_INVALID, "lstm_cell_bf16: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  float* gi = gates_ws;
  float* gh = gates_ws + (size_t)N * 4 * P;
  if (int r = linear_launch<LIN_F32>(x, w_ih, nullptr, nullptr, gi, N, K, 4 * P, st)) return r;
  if (int r = linear_launch<LIN_F32>(h, w_hh, nullptr, nullptr, gh, N, P, 4 * P, st)) return r;
  const long long total = (long long)N * P;
  const int blocks = (int)std::min<long long>((total + 255) / 256, 148ll * 16);
  lstm_postop_f32_kernel<<<blocks, 256, 0, st>>>(gi, gh, bias, c, hy, cy, N, P);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

int rnnt_greedy_decode_update(const int64_t* symbols, int32_t* symbols_added, int32_t* res, int32_t* res_idx, const uint16_t* f,
                              const int32_t* f_lens, int32_t* time_idx, uint16_t* fi, int32_t* pre_g, uint16_t* pre_hg0,
                              uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, const uint16_t* hg0, const uint16_t* hg1,
                              const float* cg0, const float* cg1, int N, int L, int C, int P, int32_t* unfinished, void* stream) {
  if (!symbols || !symbols_added || !res || !res_idx || !f || !f_lens || !time_idx || !fi || !pre_g || !pre_hg0 || !pre_hg1 ||
      !pre_cg0 || !pre_cg1 || !hg0 || !hg1 || !cg0 || !cg1 || !unfinished || N <= 0 || L <= 0 || C <= 0 || P <= 0)
    return fail(RNNT_ERR_INVALID, "greedy_decode_update: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemsetAsync(unfinished, 0, sizeof(int32_t), st));
  greedy_update_kernel<<<(N + 7) / 8, 256, 0, st>>>(reinterpret_cast<const long long*>(symbols), symbols_added, res, res_idx, f,
                                                    f_lens, time_idx, fi, pre_g, pre_hg0, pre_hg1, pre_cg0, pre_cg1, hg0, hg1,
                                                    cg0, cg1, N, L, C, P, unfinished);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

// test-only CUDA-core GEMM (cross-checks the tcgen05 kernel on the device at sizes beyond the CPU oracle)
int rnnt_gemm_i8_i32_simt(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, void* stream) {
  if (!A || !W || !C || M <= 0 || N <= 0 || K <= 0 || (K & 3)) return fail(RNNT_ERR_INVALID, "gemm_simt: bad arguments");
  if (M > 65535) return fail(RNNT_ERR_INVALID, "gemm_simt: M <= 65535 (test helper)");
  dim3 grid((N + 127) / 128, (unsigned)M);
  gemm_i8_i32_simt_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(A, W, C, M, N, K);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

}  // extern "C"
#ifdef RNNT_DEBUG_HOOKS
static int g_gemm_debug = 0;
extern "C" int rnnt_debug_set_gemm_flags(int flags) { g_gemm_debug = flags; return RNNT_OK; }
#endif
// cluster size of the projection GEMM: CTAs of a cluster share the W operand of an n tile by TMA multicast (gemm_i8.cu).
// Process-wide tuning knob (results do not depend on it): rnnt_tune_gemm_cluster / environment variable RNNT_GEMM_CLUSTER.
static int g_gemm_cluster = -1;
static int gemm_cluster_pref() {
  int& v = g_gemm_cluster;
  if (v < 0) {
    const char* s = getenv("RNNT_GEMM_CLUSTER");
    v = s ? atoi(s) : 1;        // measured on B200: 10.2 / 10.4 / 12.1 ms per bench step for 1 / 2 / 4 — the operand traffic is not the bound
    if (v != 1 && v != 2 && v != 4) v = 1;
  }
  return v;
}

// CTA-pair variant (gemm_i8_pair.cu, tcgen05 cta_group::2): rnnt_tune_gemm_pair / environment variable RNNT_GEMM_PAIR.
// On by default: 10.0 vs 10.3 ms per bench step on B200 (a third less L2 -> SM operand traffic per tile).
static int g_gemm_pair = -1;
static bool gemm_pair_pref() {
  if (g_gemm_pair < 0) {
    const char* s = getenv("RNNT_GEMM_PAIR");
    g_gemm_pair = s ? (atoi(s) != 0) : 1;
  }
  return g_gemm_pair != 0;
}

static int gemm_launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, int num_m, int num_n, int num_k,
                            int* status, const RowSkip& skip, int gb_mode, cudaStream_t st, bool* unsupported) {
  auto kern = gemm_i8_i32_pair_kernel;
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmPairSmem::TOTAL));
  const int sms = sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(GEMM_I8_THREADS);
  cfg.dynamicSmemBytes = GemmPairSmem::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int pairs = std::max(1, (sms > 0 ? sms : 148) / 2);
  cfg.gridDim = dim3(pairs * 2);
  int max_active = 0;
  if (cudaOccupancyMaxActiveClusters(&max_active, kern, &cfg) != cudaSuccess || max_active < 1) {
    cudaGetLastError();
    *unsupported = true;
    return RNNT_OK;
  }
  const int num_super = ((num_m + 1) / 2) * num_n;
  pairs = std::max(1, std::min(std::min(pairs, max_active), num_super));
  cfg.gridDim = dim3(pairs * 2);
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode);
  if (e != cudaSuccess) return fail(RNNT_ERR_CUDA, "gemm_i8_i32 (pair) launch failed: %s", cudaGetErrorString(e));
  return RNNT_OK;
}

template <int BN, int CM>
static int gemm_launch_cm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, int num_m, int num_n, int num_k,
                          int* status, const RowSkip& skip, int gb_mode, cudaStream_t st, bool* unsupported,
                          int32_t* c_direct, long long M, int ldc) {
  auto kern = gemm_i8_i32_kernel<BN, CM>;
  // the opt-in is per device/context: set on every launch (a cached per-process flag would skip it on a second GPU)
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmSmem<BN>::TOTAL));
  const int sms = sm_count();
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(GEMM_I8_THREADS);
  cfg.dynamicSmemBytes = GemmSmem<BN>::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  int clusters = std::max(1, (sms > 0 ? sms : 148) / CM);
  if (CM > 1) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CM; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    // a persistent kernel must not ask for more clusters than can be resident at once (GPC boundaries limit them)
    cfg.gridDim = dim3(clusters * CM);
    int max_active = 0;
    if (cudaOccupancyMaxActiveClusters(&max_active, kern, &cfg) != cudaSuccess || max_active < 1) {
      cudaGetLastError();
      *unsupported = true;
      return RNNT_OK;
    }
    clusters = std::min(clusters, max_active);
  }
  const int num_super = ((num_m + CM - 1) / CM) * num_n;
  clusters = std::max(1, std::min(clusters, num_super));
  cfg.gridDim = dim3(clusters * CM);
  int gb = gb_mode;
#ifdef RNNT_DEBUG_HOOKS
  gb |= g_gemm_debug << 8;
#endif
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, num_m, num_n, num_k, status, skip, gb, c_direct, M, ldc);
  if (e != cudaSuccess) return fail(RNNT_ERR_CUDA, "gemm_i8_i32 launch failed: %s", cudaGetErrorString(e));
  return RNNT_OK;
}

template <int BN>
static int gemm_launch(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, const RowSkip& skip, cudaStream_t st,
                       bool gate_blocked = false) {
  const int num_m = (int)((M + GEMM_BM - 1) / GEMM_BM), num_n = N / BN, num_k = (K + GEMM_BK - 1) / GEMM_BK;
  // clusters pay off when there are several m tiles per n tile; small problems and BN < 256 run as single CTAs
  int cm = (BN == 256 && num_m >= 8) ? gemm_cluster_pref() : 1;
  bool pair = BN == 256 && num_m >= 2 && gemm_pair_pref();
  // epilogue: staged in shared memory and written by TMA (default), or registers -> global directly (RNNT_GEMM_EPILOGUE=direct;
  // measured slower on B200: 11.3 vs 10.2 ms per bench step)
  static const bool direct_epilogue = [] { const char* e = getenv("RNNT_GEMM_EPILOGUE"); return e && !strcmp(e, "direct"); }();
  int32_t* c_direct = direct_epilogue ? C : nullptr;
  int* status = device_status();
  if (!status) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  for (;;) {
    CUtensorMap ta, tb, tc;
    int gb_mode = 0;              // 0 row-major result, 1 gate-blocked as 32-byte rows, 2 gate-blocked as 1 KB runs
    {
      uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[1] = {(uint64_t)K};
      uint32_t box[2] = {GEMM_BK, GEMM_BM};
      if (int r = make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, A, dims, str, box)) return r;
    }
    {
      uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[1] = {(uint64_t)K};
      uint32_t box[2] = {GEMM_BK, (uint32_t)(pair ? BN / 2 : BN / cm)};
      if (int r = make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, W, dims, str, box)) return r;
    }
    if (!gate_blocked) {
      uint64_t dims[2] = {(uint64_t)N, (uint64_t)M}, str[1] = {(uint64_t)N * 4};
      uint32_t box[2] = {32, 32};
      if (int r = make_tmap(&tc, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, C, dims, str, box)) return r;
    } else if (M % 32 == 0) {
      // gate-blocked projection for the recurrent kernel: C[slice 32][gate 4][unit octet 4][row M][8] int32, column
      // gate*1024 + slice*32 + octet*8 + u of the row-major result.  A 32-column chunk of the epilogue is one (gate, slice),
      // and its 32 rows x 8 units of one octet are one contiguous KB: box = 256 int32 x 1 run x 4 octets.
      uint64_t dims[5] = {256, (uint64_t)M / 32, 4, 4, 32};
      uint64_t str[4] = {1024, (uint64_t)M * 32, (uint64_t)M * 128, (uint64_t)M * 512};
      uint32_t box[5] = {256, 1, 4, 1, 1};
      if (int r = make_tmap(&tc, CU_TENSOR_MAP_DATA_TYPE_INT32, 5, C, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return r;
      gb_mode = 2;
    } else {
      // the same layout for any M, as rows of 32 bytes: box = 8 units x 32 rows x 4 octets
      uint64_t dims[5] = {8, (uint64_t)M, 4, 4, 32};
      uint64_t str[4] = {32, (uint64_t)M * 32, (uint64_t)M * 128, (uint64_t)M * 512};
      uint32_t box[5] = {8, 32, 4, 1, 1};
      if (int r = make_tmap(&tc, CU_TENSOR_MAP_DATA_TYPE_INT32, 5, C, dims, str, box, CU_TENSOR_MAP_SWIZZLE_32B)) return r;
      gb_mode = 1;
    }
    bool unsupported = false;
    int r;
    if (pair) {
      r = gemm_launch_pair(ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode, st, &unsupported);
      if (r) return r;
      if (!unsupported) return RNNT_OK;
      pair = false;                                // no resident CTA pair possible here: the single-CTA kernel
      continue;
    }
    if constexpr (BN == 256) {
      if (cm == 4) r = gemm_launch_cm<256, 4>(ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode, st, &unsupported, c_direct, M, N);
      else if (cm == 2) r = gemm_launch_cm<256, 2>(ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode, st, &unsupported, c_direct, M, N);
      else r = gemm_launch_cm<256, 1>(ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode, st, &unsupported, c_direct, M, N);
    } else {
      r = gemm_launch_cm<BN, 1>(ta, tb, tc, num_m, num_n, num_k, status, skip, gb_mode, st, &unsupported, c_direct, M, N);
    }
    if (r) return r;
    if (!unsupported) return RNNT_OK;
    cm >>= 1;                                    // cluster launch not possible here: halve the cluster
    if (cm < 1) return fail(RNNT_ERR_CUDA, "gemm_i8_i32: no launchable configuration");
  }
}

extern "C" {
int rnnt_gemm_i8_i32(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, void* stream) {
  return rnnt_gemm_i8_i32_ragged(A, W, C, M, N, K, 0, nullptr, stream);
}
int rnnt_gemm_i8_i32_ragged(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, int rows_per_step,
                            const int32_t* tile_steps_host, void* stream) {
  if (!A || !W || !C || M < 0 || N <= 0 || K <= 0)
    return fail(RNNT_ERR_INVALID, "gemm_i8_i32: null pointer or non-positive size");
  if ((K & 15) || (N & 63) || !aligned(A, 16) || !aligned(W, 16) || !aligned(C, 16))
    return fail(RNNT_ERR_INVALID, "gemm_i8_i32: need K %% 16 == 0, N %% 64 == 0, 16-byte aligned pointers (M=%lld N=%d K=%d)",
                (long long)M, N, K);
  if (M == 0) return RNNT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  RowSkip skip;
  if (int r = make_skip(&skip, M, rows_per_step, tile_steps_host, GEMM_BM)) return r;
  if (N % 256 == 0) return gemm_launch<256>(A, W, C, M, N, K, skip, st);
  if (N % 128 == 0) return gemm_launch<128>(A, W, C, M, N, K, skip, st);
  return gemm_launch<64>(A, W, C, M, N, K, skip, st);
}

int rnnt_gemm_i8_i32_gate_blocked_ragged(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, int rows_per_step,
                                         const int32_t* tile_steps_host, void* stream) {
  if (!A || !W || !C || M < 0 || K <= 0) return fail(RNNT_ERR_INVALID, "gemm_i8_i32_gate_blocked: null pointer or non-positive size");
  if (N != 4 * REC_H) return fail(RNNT_ERR_INVALID, "gemm_i8_i32_gate_blocked: N must be 4096 (4 gates x 1024 units), got %d", N);
  if ((K & 15) || !aligned(A, 16) || !aligned(W, 16) || !aligned(C, 32))
    return fail(RNNT_ERR_INVALID, "gemm_i8_i32_gate_blocked: need K %% 16 == 0, 16-byte aligned A/W, 32-byte aligned C");
  if (M == 0) return RNNT_OK;
  RowSkip skip;
  if (int r = make_skip(&skip, M, rows_per_step, tile_steps_host, GEMM_BM)) return r;
  return gemm_launch<256>(A, W, C, M, N, K, skip, (cudaStream_t)stream, true);
}

}  // extern "C"

template <int BN>
static int gemm_bf16_launch(const uint16_t* A, const uint16_t* W, uint16_t* C, int64_t M, int N, int K, const RowSkip& skip,
                            cudaStream_t st) {
  using S = GemmBf16Smem<BN>;
  CUtensorMap ta, tb, tc;
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[1] = {(uint64_t)K * 2};
    uint32_t box[2] = {(uint32_t)S::BK_ELEMS, GEMM_BM};
    if (int r = make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, dims, str, box)) return r;
  }
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[1] = {(uint64_t)K * 2};
    uint32_t box[2] = {(uint32_t)S::BK_ELEMS, (uint32_t)BN};
    if (int r = make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, W, dims, str, box)) return r;
  }
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M}, str[1] = {(uint64_t)N * 2};
    uint32_t box[2] = {32, 32};
    if (int r = make_tmap(&tc, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, C, dims, str, box, CU_TENSOR_MAP_SWIZZLE_64B)) return r;
  }
  int* status = device_status();
  if (!status) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  auto kern = gemm_bf16_bf16_kernel<BN>;
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
  const int num_m = (int)((M + GEMM_BM - 1) / GEMM_BM), num_n = N / BN, num_k = (K + S::BK_ELEMS - 1) / S::BK_ELEMS;
  const int sms = sm_count();
  const int grid = std::max(1, std::min(sms > 0 ? sms : 148, num_m * num_n));
  kern<<<grid, GEMM_THREADS, S::TOTAL, st>>>(ta, tb, tc, num_m, num_n, num_k, status, skip);
  CUDA_TRY(cudaGetLastError());
  return RNNT_OK;
}

extern "C" {
int rnnt_gemm_bf16_bf16_ragged(const uint16_t* A, const uint16_t* W, uint16_t* C, int64_t M, int N, int K, int rows_per_step,
                               const int32_t* tile_steps_host, void* stream) {
  if (!A || !W || !C || M < 0 || N <= 0 || K <= 0) return fail(RNNT_ERR_INVALID, "gemm_bf16_bf16: null pointer or non-positive size");
  if ((K & 7) || (N & 63) || !aligned(A, 16) || !aligned(W, 16) || !aligned(C, 16))
    return fail(RNNT_ERR_INVALID, "gemm_bf16_bf16: need K %% 8 == 0, N %% 64 == 0, 16-byte aligned pointers (M=%lld N=%d K=%d)",
                (long long)M, N, K);
  if (M == 0) return RNNT_OK;
  RowSkip skip;
  if (int r = make_skip(&skip, M, rows_per_step, tile_steps_host, GEMM_BM)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  if (N % 256 == 0) return gemm_bf16_launch<256>(A, W, C, M, N, K, skip, st);
  if (N % 128 == 0) return gemm_bf16_launch<128>(A, W, C, M, N, K, skip, st);
  return gemm_bf16_launch<64>(A, W, C, M, N, K, skip, st);
}
int rnnt_gemm_bf16_bf16(const uint16_t* A, const uint16_t* W, uint16_t* C, int64_t M, int N, int K, void* stream) {
  return rnnt_gemm_bf16_bf16_ragged(A, W, C, M, N, K, 0, nullptr, stream);
}

#ifdef RNNT_DEBUG_HOOKS
// debug hook: device buffer [T][16] int64 receiving clock64 stamps of CTA (0,0) of the next recurrent launches (NULL = off)
static long long* g_rec_trace = nullptr;
int rnnt_debug_set_recurrent_trace(void* device_buffer) { g_rec_trace = reinterpret_cast<long long*>(device_buffer); return RNNT_OK; }
#endif

// ------------------------------------------------------------------------------------------------
// recurrent kernel.  workspace: [0,1024) reserved | flags [num_btiles][32] int | h ping-pong [2][N_pad][1024] int8 (K-chunk-major)
//                             | c scratch [N_pad][1024] binary16 (blocked)
static int64_t rec_flags_bytes(int N) { return round_up((int64_t)((N + REC_BM - 1) / REC_BM) * REC_FLAG_STRIDE * 4, 256); }
int64_t rnnt_lstm_recurrent_workspace_bytes(int N) {
  if (N <= 0) return 0;
  const int64_t n_pad = round_up(N, REC_BM);
  return 1024 + rec_flags_bytes(N) + 2 * n_pad * REC_H + 1024 + n_pad * REC_H * 2;
}

}  // extern "C"

// Launch with `cs` slice-CTAs per thread-block cluster (h tiles are TMA-multicast inside a cluster).  Cooperative so that
// all CTAs are co-resident (they exchange h through per-tile counters).  Returns cudaErrorCooperativeLaunchTooLarge /
// cluster errors to the caller, which falls back to a smaller cluster.
template <int CS, int HS>
static cudaError_t rec_launch_cs(const CUtensorMap& tw, const CUtensorMap& th0, const CUtensorMap& th,
                                 const RecParams& p, int num_slices, int grid_y, cudaStream_t st) {
  auto kern = lstm_recurrent_kernel<CS, HS>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, RecSmemT<HS>::TOTAL);   // per device
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_slices, grid_y);
  cfg.blockDim = dim3(REC_THREADS);
  cfg.dynamicSmemBytes = RecSmemT<HS>::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  attrs[na].id = cudaLaunchAttributeCooperative; attrs[na].val.cooperative = 1; ++na;
  if (CS > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = CS; attrs[na].val.clusterDim.y = 1; attrs[na].val.clusterDim.z = 1; ++na;
  }
  cfg.attrs = attrs; cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, tw, th0, th, p);
}
template <int HS>
static int rec_launch_hs(const CUtensorMap& tw, const CUtensorMap& th0, const CUtensorMap& th,
                         const RecParams& p, int num_slices, int grid_y, int cs_max, cudaStream_t st) {
  cudaError_t e = cudaErrorUnknown;
  if (cs_max >= 8) { e = rec_launch_cs<8, HS>(tw, th0, th, p, num_slices, grid_y, st); if (e == cudaSuccess) return RNNT_OK; cudaGetLastError(); }
  if (cs_max >= 4) { e = rec_launch_cs<4, HS>(tw, th0, th, p, num_slices, grid_y, st); if (e == cudaSuccess) return RNNT_OK; cudaGetLastError(); }
  if (cs_max >= 2) { e = rec_launch_cs<2, HS>(tw, th0, th, p, num_slices, grid_y, st); if (e == cudaSuccess) return RNNT_OK; cudaGetLastError(); }
  e = rec_launch_cs<1, HS>(tw, th0, th, p, num_slices, grid_y, st);
  if (e != cudaSuccess) return fail(RNNT_ERR_CUDA, "lstm_recurrent launch failed: %s", cudaGetErrorString(e));
  return RNNT_OK;
}
static int rec_launch(const CUtensorMap& tw, const CUtensorMap& th0, const CUtensorMap& th,
                      const RecParams& p, int hs, int grid_y, int cs_max, cudaStream_t st) {
  return hs == 16 ? rec_launch_hs<16>(tw, th0, th, p, REC_H / 16, grid_y, cs_max, st)
                  : rec_launch_hs<32>(tw, th0, th, p, REC_H / 32, grid_y, cs_max, st);
}

static int g_rec_cluster = -1;
static int rec_cluster_pref() {
  int& v = g_rec_cluster;
  if (v < 0) {
    const char* s = getenv("RNNT_REC_CLUSTER");   // 1, 2 or 4 slice-CTAs per cluster (tuning knob; default 1: L2 already
    v = s ? atoi(s) : 1;                          // de-duplicates <= 4 concurrent readers of a line, measured no gain)
    if (v != 1 && v != 2 && v != 4 && v != 8) v = 1;
  }
  return v;
}

extern "C" {
#ifdef RNNT_DEBUG_HOOKS
static int g_rec_debug = 0;
int rnnt_debug_set_recurrent_flags(int flags) { g_rec_debug = flags; return RNNT_OK; }
#endif
int rnnt_tune_gemm_pair(int on) {
  g_gemm_pair = on ? 1 : 0;
  return RNNT_OK;
}
int rnnt_tune_gemm_cluster(int ctas) {
  if (ctas != 1 && ctas != 2 && ctas != 4) return fail(RNNT_ERR_INVALID, "gemm cluster size must be 1, 2 or 4");
  g_gemm_cluster = ctas;
  return RNNT_OK;
}
int rnnt_tune_recurrent_cluster(int ctas) {
  if (ctas != 1 && ctas != 2 && ctas != 4 && ctas != 8) return fail(RNNT_ERR_INVALID, "recurrent cluster size must be 1, 2, 4 or 8");
  g_rec_cluster = ctas;
  return RNNT_OK;
}
static int rec_run(const int32_t* G, int g_blocked, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                   float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                   const int32_t* tile_steps_host, void* workspace, void* stream, int* status_word = nullptr);
int rnnt_lstm_recurrent_i8_ragged(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                                  float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                                  const int32_t* tile_steps_host, void* workspace, void* stream) {
  return rec_run(G, 0, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, hx, cx, y, T, N, H, tile_steps_host, workspace, stream);
}
int rnnt_lstm_recurrent_i8_gate_blocked_ragged(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale,
                                               float in_scale, float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y,
                                               int T, int N, int H, const int32_t* tile_steps_host, void* workspace, void* stream) {
  return rec_run(G, 1, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, hx, cx, y, T, N, H, tile_steps_host, workspace, stream);
}
static int rec_run(const int32_t* G, int g_blocked, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                   float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                   const int32_t* tile_steps_host, void* workspace, void* stream, int* status_word) {
  if (!G || !w_hh || !bias || !hx || !cx || !y || !workspace) return fail(RNNT_ERR_INVALID, "lstm_recurrent: null pointer");
  if (H != REC_H) return fail(RNNT_ERR_INVALID, "lstm_recurrent: hidden size must be 1024 (got %d)", H);
  if (T < 0 || N <= 0) return fail(RNNT_ERR_INVALID, "lstm_recurrent: need T >= 0, N > 0 (T=%d N=%d)", T, N);
  if (!aligned(G, 32) || !aligned(w_hh, 16) || !aligned(workspace, 256) || !aligned(y, 8) || !aligned(cx, 16) || !aligned(hx, 16))
    return fail(RNNT_ERR_INVALID, "lstm_recurrent: misaligned buffer");
  if (T == 0) return RNNT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int num_btiles = (N + REC_BM - 1) / REC_BM;
  const int64_t n_pad = (int64_t)num_btiles * REC_BM;
  if (num_btiles > REC_MAX_TILES) return fail(RNNT_ERR_INVALID, "lstm_recurrent: N <= %d per call (got %d)", REC_MAX_TILES * REC_BM, N);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  int* flags = reinterpret_cast<int*>(ws + 1024);
  int8_t* h_buf = reinterpret_cast<int8_t*>(ws + 1024 + rec_flags_bytes(N));
  h_buf = reinterpret_cast<int8_t*>(round_up((int64_t)reinterpret_cast<uintptr_t>(h_buf), 1024));
  int* status = status_word ? status_word : device_status();     // status_word: only the static-scratch entry point passes one
  if (!status) return fail(RNNT_ERR_CUDA, "no usable CUDA device");

  RecParams p;
  bool any = false;
  for (int b = 0; b < REC_MAX_TILES; ++b) {
    int s = 0;
    if (b < num_btiles) {
      s = tile_steps_host ? tile_steps_host[b] : T;
      if (s < 0 || s > T) return fail(RNNT_ERR_INVALID, "lstm_recurrent: tile_steps[%d] = %d outside [0, T=%d]", b, s, T);
    }
    p.tile_steps[b] = s;
    any |= s > 0;
  }
  if (!any) return RNNT_OK;

  CUDA_TRY(cudaMemsetAsync(flags, 0, rec_flags_bytes(N), st));
  uint16_t* c_ws = reinterpret_cast<uint16_t*>(h_buf + 2 * n_pad * REC_H);

  // Batches of <= 4 tiles run ONE dependency chain per tile group and timestep: there the kernel with 16-unit slices
  // (64 slice CTAs x 2 tile groups, 7 ring slots, 4 units per epilogue thread) has the shorter chain (kernels.cuh).
  static const bool hs16_ok = [] { const char* e = getenv("RNNT_REC_HS16"); return e ? atoi(e) != 0 : true; }();
  const int hs = (!status_word && hs16_ok && num_btiles <= 4) ? 16 : REC_HS;
  const int max_groups = hs == 16 ? 2 : REC_GRID_Y;
  const int grid_y = num_btiles < max_groups ? num_btiles : max_groups;
  CUtensorMap tw, th0, th;
  {   // W_hh [4 gates][1024 units][1024 k]: box = 128 B of k x 32 units x 4 gates
    uint64_t dims[3] = {REC_H, REC_H, 4}, str[2] = {REC_H, (uint64_t)REC_H * REC_H};
    uint32_t box[3] = {128, (uint32_t)hs, 4};
    if (int r = make_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w_hh, dims, str, box)) return r;
  }
  {   // h_0 = the caller's hx [N][1024] (rows >= N read as zero)
    uint64_t dims[2] = {REC_H, (uint64_t)N}, str[1] = {REC_H};
    uint32_t box[2] = {128, REC_BM};
    if (int r = make_tmap(&th0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, hx, dims, str, box)) return r;
  }
  {   // h ping-pong, K-chunk-major: rows of 128 B, one 128-row chunk = 16 KB contiguous, stored 128-byte-swizzled by the
      // kernel itself (read back with one bulk copy per chunk; this map, without swizzle, serves the multicast variants)
    uint64_t dims[2] = {128, (uint64_t)(2 * n_pad * REC_KC)}, str[1] = {128};
    uint32_t box[2] = {128, REC_BM};
    if (int r = make_tmap(&th, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, h_buf, dims, str, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return r;
  }
  // ---- deal the batch tiles to the grid's tile groups.  Cost model from the traces (profiles/r02): a work item costs a CTA
  // ~8K cycles of throughput, but one (tile, step) dependency chain is ~16K cycles long, so a group with 1 or 2 tiles in
  // flight advances one step per ~16K cycles and only the third, fourth ... tile adds time: in item units
  // cost(group) = 2 * longest + sum of all but the two longest.  Dealing tiles round-robin leaves the group that owns
  // the longest tile with long companions too; instead the tiles go, longest first, to the group that keeps the maximum
  // group cost lowest (the longest tile's group then takes the second longest for free and short ones after that).
  {
    int order[REC_MAX_TILES];
    for (int b = 0; b < num_btiles; ++b) order[b] = b;
    std::stable_sort(order, order + num_btiles, [&](int a, int b) { return p.tile_steps[a] > p.tile_steps[b]; });
    for (int g = 0; g < REC_GRID_Y; ++g) { p.group_count[g] = 0; for (int i = 0; i < REC_MAX_TILES; ++i) p.group_tile[g][i] = 0; }
    auto cost = [&](int g, int extra) {          // tiles arrive longest first, so a group's list is sorted
      long long c = 0;
      const int n = p.group_count[g] + (extra >= 0 ? 1 : 0);
      for (int i = 0; i < n; ++i) {
        const int s = i < p.group_count[g] ? p.tile_steps[p.group_tile[g][i]] : extra;
        c += (i == 0) ? 2ll * s : (i == 1 ? 0 : s);
      }
      return c;
    };
    for (int k = 0; k < num_btiles; ++k) {
      const int b = order[k];
      int best = 0;
      if (num_btiles <= grid_y) {
        best = k;                                  // one tile per group while there are groups to spare
      } else {
        long long best_max = -1, best_sum = -1;
        for (int g = 0; g < grid_y; ++g) {
          long long mx = 0, sum = 0;
          for (int j = 0; j < grid_y; ++j) { const long long c = cost(j, j == g ? p.tile_steps[b] : -1); mx = std::max(mx, c); sum += c; }
          if (best_max < 0 || mx < best_max || (mx == best_max && sum < best_sum)) { best_max = mx; best_sum = sum; best = g; }
        }
      }
      p.group_tile[best][p.group_count[best]++] = b;
    }
  }
  int used_groups = 0;                              // groups are filled in index order: launch only the non-empty ones
  while (used_groups < grid_y && p.group_count[used_groups] > 0) ++used_groups;
  p.g = G; p.g_blocked = g_blocked; p.bias = bias; p.h_buf = h_buf; p.hx = hx; p.cx = cx; p.c_ws = c_ws; p.y = y; p.flags = flags; p.status = status;
  p.rb_scale = rb_scale; p.in_scale = in_scale; p.out_scale = out_scale;
  p.T = T; p.N = N; p.num_btiles = num_btiles; p.skip_quant_y = skip_quant_y ? 1 : 0;
#ifdef RNNT_DEBUG_HOOKS
  p.trace = g_rec_trace;
  p.debug = g_rec_debug;
#else
  p.trace = nullptr;
  p.debug = 0;
#endif
  if (status_word) {      // static-scratch entry point: no library state is created; overlapping calls race by contract
    if (int r = rec_launch(tw, th0, th, p, hs, used_groups, rec_cluster_pref(), st)) return r;
  } else {
    CoopOrder order;
    if (int r = order.begin(st)) return r;
    const int r = rec_launch(tw, th0, th, p, hs, used_groups, rec_cluster_pref(), st);
    const int r2 = order.end();
    if (r) return r;
    if (r2) return r2;
  }
  // tiles that ran a single step left their final h in the ping-pong buffer (see lstm_recurrent.cu): move it to hx
  unsigned one_step = 0;
  for (int b = 0; b < num_btiles; ++b) one_step |= (p.tile_steps[b] == 1) ? (1u << b) : 0u;
  if (one_step) {
    rec_copy_h1_kernel<<<num_btiles, 256, 0, st>>>(h_buf, hx, num_btiles, N, one_step);
    CUDA_TRY(cudaGetLastError());
  }
  return RNNT_OK;
}

int rnnt_lstm_recurrent_i8(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                           float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                           void* workspace, void* stream) {
  return rnnt_lstm_recurrent_i8_ragged(G, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, hx, cx, y, T, N, H, nullptr,
                                       workspace, stream);
}

// ---- rnnt_lstm_recurrent_i8_static_scratch: the recurrent kernel for callers that cannot allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  Its scratch — the device status word, the per-tile release
// counters, the h ping-pong buffer and the blocked c scratch — lives in ONE file-scope device array of fixed size; every
// GPU that loads the library holds these bytes until the process exits.  All calls share the array: it keeps the last
// call's contents (the status word and the counters are cleared at the start of every call), and two calls that overlap
// on one GPU — different streams, or graph branches — race on it.
constexpr int REC_STATIC_MAX_N = 512;                       // batch rows per call
constexpr int64_t REC_STATIC_STATUS_BYTES = 256;            // the status word, in front of the workspace proper
constexpr int64_t REC_STATIC_WS_BYTES =                     // = rnnt_lstm_recurrent_workspace_bytes(REC_STATIC_MAX_N)
    1024 + (REC_STATIC_MAX_N / REC_BM) * REC_FLAG_STRIDE * 4 + 2ll * REC_STATIC_MAX_N * REC_H + 1024 + 2ll * REC_STATIC_MAX_N * REC_H;
static_assert(REC_STATIC_MAX_N % REC_BM == 0 && ((REC_STATIC_MAX_N / REC_BM) * REC_FLAG_STRIDE * 4) % 256 == 0, "flag block is a multiple of 256 B");
static __device__ __align__(256) unsigned char g_rec_static_scratch[REC_STATIC_STATUS_BYTES + REC_STATIC_WS_BYTES];   // 2,099,968 B

// Device address of the array on the current device.  In host code `g_rec_static_scratch` is only a handle for the
// runtime, not an address a kernel could use: cudaGetSymbolAddress translates it, once per device, with that device current.
static int rec_static_scratch_ptr(uint8_t** out) {
  static uint8_t* ptr[64] = {nullptr};
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* q = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&q, g_rec_static_scratch);
    if (e != cudaSuccess || !q)
      return fail(RNNT_ERR_CUDA, "cudaGetSymbolAddress(g_rec_static_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    ptr[dev] = reinterpret_cast<uint8_t*>(q);
  }
  *out = ptr[dev];
  return RNNT_OK;
}

int rnnt_lstm_recurrent_i8_static_scratch(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                                          float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                                          int32_t* status_out_host, void* stream) {
  if (N > REC_STATIC_MAX_N)
    return fail(RNNT_ERR_INVALID, "lstm_recurrent_static_scratch: N = %d exceeds the static scratch (N <= %d per call); use "
                "rnnt_lstm_recurrent_i8 with a caller-owned workspace", N, REC_STATIC_MAX_N);
  if (N > 0 && rnnt_lstm_recurrent_workspace_bytes(N) > REC_STATIC_WS_BYTES)
    return fail(RNNT_ERR_INVALID, "lstm_recurrent_static_scratch: workspace for N = %d does not fit the static scratch", N);
  uint8_t* scratch = nullptr;
  if (int r = rec_static_scratch_ptr(&scratch)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  int* status = reinterpret_cast<int*>(scratch);
  // must start at zero: the status word (kernels bail out while it is set); rec_run clears the release counters itself
  CUDA_TRY(cudaMemsetAsync(status, 0, REC_STATIC_STATUS_BYTES, st));
  if (int r = rec_run(G, 0, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, hx, cx, y, T, N, H, nullptr,
                      scratch + REC_STATIC_STATUS_BYTES, stream, status))
    return r;
  if (status_out_host) CUDA_TRY(cudaMemcpyAsync(status_out_host, status, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  return RNNT_OK;
}

int rnnt_lstm_layer_i8(const int8_t* x, int8_t* hx, uint16_t* cx, const int8_t* w_ih, const int8_t* w_hh, const float* bias,
                       float rb_scale, float in_scale, float out_scale, int skip_quant_y, void* y, int T, int N, int K, int H,
                       int32_t* g_ws, void* workspace, void* stream) {
  if (!g_ws) return fail(RNNT_ERR_INVALID, "lstm_layer: null projection workspace");
  if (T == 0) return RNNT_OK;
  if (int r = rnnt_gemm_i8_i32(x, w_ih, g_ws, (int64_t)T * N, 4 * H, K, stream)) return r;
  return rnnt_lstm_recurrent_i8(g_ws, w_hh, bias, rb_scale, in_scale, out_scale, skip_quant_y, hx, cx, y, T, N, H, workspace,
                                stream);
}

// ------------------------------------------------------------------------------------------------
// greedy decode.  workspace: [0,256) global_sync | group_cnt [8][32] int | lane_state [N][8] int | z [N][512] f32 |
//                            hg1 [2][N][320] bf16 | cg1 [2][N][320] f32
#ifdef RNNT_DEBUG_HOOKS
static long long* g_dec_trace = nullptr;
static int g_dec_trace_cap = 0, g_dec_trace_block = 0, g_dec_skew = 0;
int rnnt_debug_set_decode_trace_block(int block) { g_dec_trace_block = block; return RNNT_OK; }
int rnnt_debug_set_decode_skew(int cycles) { g_dec_skew = cycles; return RNNT_OK; }
int rnnt_debug_set_decode_trace(void* device_buffer, int max_iters) {
  g_dec_trace = reinterpret_cast<long long*>(device_buffer); g_dec_trace_cap = max_iters; return RNNT_OK;
}
#endif
int64_t rnnt_greedy_decode_workspace_bytes(int N) {
  if (N <= 0) return 0;
  return 256 + 1024 + round_up((int64_t)N * 32, 256) + round_up((int64_t)N * DEC_J * 4, 256) +
         round_up((int64_t)2 * N * DEC_P * 2, 256) + round_up((int64_t)2 * N * DEC_P * 4, 256);
}

static int decode_run(bool tensor_cores, const uint16_t* y1, const int32_t* f_lens, int T2, int N, const rnnt_decode_weights* w,
                      int32_t* res, int L, int32_t* res_idx, int32_t* symbols_added, int32_t* time_idx, int32_t* pre_g,
                      uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, int32_t* iters_out, void* workspace,
                      void* stream) {
  if (!y1 || !f_lens || !w || !res || !res_idx || !symbols_added || !time_idx || !pre_g || !pre_hg0 || !pre_hg1 ||
      !pre_cg0 || !pre_cg1 || !workspace)
    return fail(RNNT_ERR_INVALID, "greedy_decode: null pointer");
  if (!w->embed || !w->w_ih[0] || !w->w_ih[1] || !w->w_hh[0] || !w->w_hh[1] || !w->bias[0] || !w->bias[1] || !w->w_pred ||
      !w->b1 || !w->w2 || !w->b2)
    return fail(RNNT_ERR_INVALID, "greedy_decode: null weight pointer");
  if (T2 <= 0 || N <= 0 || L <= 0) return fail(RNNT_ERR_INVALID, "greedy_decode: need T2, N, L > 0 (T2=%d N=%d L=%d)", T2, N, L);
  if (!aligned(workspace, 256) || !aligned(y1, 16) || !aligned(w->b1, 16) || !aligned(w->w_pred, 8) || !aligned(w->w2, 16) || !aligned(w->embed, 16) || !aligned(pre_hg0, 16) || !aligned(pre_hg1, 16) || !aligned(w->w_ih[0], 4) ||
      !aligned(w->w_hh[0], 4) || !aligned(w->w_ih[1], 4) || !aligned(w->w_hh[1], 4))
    return fail(RNNT_ERR_INVALID, "greedy_decode: misaligned buffer");
  if (tensor_cores && (!aligned(w->w_ih[0], 16) || !aligned(w->w_hh[0], 16) || !aligned(w->w_ih[1], 16) || !aligned(w->w_hh[1], 16) ||
                       !aligned(w->w_pred, 16)))
    return fail(RNNT_ERR_INVALID, "greedy_decode: prediction / joint weights must be 16-byte aligned");
  const int NG = 7;
  const int LG = (N + NG - 1) / NG;
  if (LG > DEC_MAXLG) return fail(RNNT_ERR_INVALID, "greedy_decode: N <= %d per call (got %d)", NG * DEC_MAXLG, N);
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  int* status = device_status();
  if (!status) return fail(RNNT_ERR_CUDA, "no usable CUDA device");
  DecParams p;
  p.embed = w->embed;
  for (int l = 0; l < 2; ++l) { p.w_ih[l] = w->w_ih[l]; p.w_hh[l] = w->w_hh[l]; p.bias[l] = w->bias[l]; }
  p.w_pred = w->w_pred; p.b1 = w->b1; p.w2 = w->w2; p.b2 = w->b2;
  p.y1 = y1; p.f_lens = f_lens; p.T2 = T2; p.N = N;
  p.res = res; p.L = L; p.res_idx = res_idx; p.symbols_added = symbols_added; p.time_idx = time_idx; p.pre_g = pre_g;
  int64_t off = 0;
  p.global_sync = reinterpret_cast<int32_t*>(ws + off); off += 256;
  p.group_cnt = reinterpret_cast<int32_t*>(ws + off); off += 1024;
  p.lane_state = reinterpret_cast<int32_t*>(ws + off); off += round_up((int64_t)N * 32, 256);
  p.z = reinterpret_cast<float*>(ws + off); off += round_up((int64_t)N * DEC_J * 4, 256);
  uint16_t* hg1 = reinterpret_cast<uint16_t*>(ws + off); off += round_up((int64_t)2 * N * DEC_P * 2, 256);
  float* cg1 = reinterpret_cast<float*>(ws + off);
  p.hg[0][0] = pre_hg0; p.hg[0][1] = pre_hg1; p.hg[1][0] = hg1; p.hg[1][1] = hg1 + (size_t)N * DEC_P;
  p.cg[0][0] = pre_cg0; p.cg[0][1] = pre_cg1; p.cg[1][0] = cg1; p.cg[1][1] = cg1 + (size_t)N * DEC_P;
  p.iters_out = iters_out; p.status = status; p.NG = NG; p.LG = LG;
  static const int max_run = [] { const char* e = getenv("RNNT_DEC_MAX_RUN"); const int v = e ? atoi(e) : 4; return v > 0 ? v : 4; }();   // measured: 2048-lane batch 11.3 / 11.6 / 12.5 / 15.0 ms for 4 / 8 / 16 / 64
  p.max_run = max_run;      // tuning knob (results do not depend on it)
#ifdef RNNT_DEBUG_HOOKS
  p.trace = g_dec_trace; p.trace_cap = g_dec_trace_cap; p.trace_block = g_dec_trace_block; p.skew = g_dec_skew;
#else
  p.trace = nullptr; p.trace_cap = 0; p.trace_block = 0; p.skew = 0;
#endif
  CUDA_TRY(cudaMemsetAsync(ws, 0, 256 + 1024, st));
  const void* kern = tensor_cores ? (const void*)greedy_decode_kernel<true> : (const void*)greedy_decode_kernel<false>;
  const int smem_bytes = tensor_cores ? DecSmemTC::TOTAL : DecSmem::TOTAL;
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));   // per device
  void* args[1] = {(void*)&p};
  CoopOrder order;
  if (int r = order.begin(st)) return r;
  const cudaError_t le = cudaLaunchCooperativeKernel(kern, dim3(NG * DEC_S), dim3(DEC_THREADS), args, smem_bytes, st);
  const int r2 = order.end();
  if (le != cudaSuccess) return fail(RNNT_ERR_CUDA, "greedy_decode launch failed: %s", cudaGetErrorString(le));
  if (r2) return r2;
  return RNNT_OK;
}

int rnnt_greedy_decode(const uint16_t* y1, const int32_t* f_lens, int T2, int N, const rnnt_decode_weights* w, int32_t* res,
                       int L, int32_t* res_idx, int32_t* symbols_added, int32_t* time_idx, int32_t* pre_g, uint16_t* pre_hg0,
                       uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, int32_t* iters_out, void* workspace, void* stream) {
  return decode_run(true, y1, f_lens, T2, N, w, res, L, res_idx, symbols_added, time_idx, pre_g, pre_hg0, pre_hg1, pre_cg0, pre_cg1,
                    iters_out, workspace, stream);
}
int rnnt_greedy_decode_exact_order(const uint16_t* y1, const int32_t* f_lens, int T2, int N, const rnnt_decode_weights* w,
                                   int32_t* res, int L, int32_t* res_idx, int32_t* symbols_added, int32_t* time_idx, int32_t* pre_g,
                                   uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, int32_t* iters_out,
                                   void* workspace, void* stream) {
  return decode_run(false, y1, f_lens, T2, N, w, res, L, res_idx, symbols_added, time_idx, pre_g, pre_hg0, pre_hg1, pre_cg0,
                    pre_cg1, iters_out, workspace, stream);
}

}  // extern "C"
