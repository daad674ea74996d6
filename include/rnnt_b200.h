/* rnnt_b200.h — C ABI of librnnt_b200.so: the sm_100a CUDA implementation of the MLPerf RNN-T int8
 * inference hot path of dbyoung18/rnnt-inference.
 *
 * Every entry point replaces one operator of the reference's plugin library
 * (TORCH_LIBRARY(intel_mlperf), mlperf_plugins/csrc/init.cpp:20-144) or one fused stretch of its
 * Python driver; the reference interface each one stands in for is cited beside it.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - the caller owns every buffer, including workspaces; nothing is allocated or freed here;
 *   - return value 0 = success, otherwise a negative code; rnnt_last_error() gives the message
 *     (thread-local).  No exception crosses this boundary (the reference throws std::runtime_error,
 *     e.g. tpps/i_linear_tpp.cpp:23,104,109, stack_time.cpp:32 — the Python wrapper re-raises
 *     RuntimeError from the code);
 *   - weights are plain row-major [out_features][in_features] (torch layout); the x86 VNNI tiling
 *     of models/quant_modules.py:177-208 is an AMX detail and is not part of this ABI;
 *   - binary16 / bfloat16 buffers are typed uint16_t.
 */
#ifndef RNNT_B200_H_
#define RNNT_B200_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RNNT_OK 0
#define RNNT_ERR_INVALID (-1)   /* bad shape / alignment / null pointer                 */
#define RNNT_ERR_CUDA (-2)      /* a CUDA runtime/driver call failed                    */
#define RNNT_ERR_DEVICE (-3)    /* device-side watchdog or protocol error (see message) */

const char* rnnt_last_error(void);
/* library/ABI version, and the SM count / arch the library sees (-1 without a usable GPU) */
int rnnt_abi_version(void);
int rnnt_device_sm_count(void);
/* Kernels run asynchronously on `stream`; their device-side watchdogs (a protocol error ends the kernel
 * instead of hanging the GPU) record a code in a status block owned by the library (first word; the wait that fired
 * first also leaves its site, CTA and epoch, which the error message reports).  This call synchronises `stream`, returns
 * RNNT_ERR_DEVICE if a code was recorded, and clears it.
 * Streams: every entry point may be called on any stream.  The two cooperative kernels (rnnt_lstm_recurrent_i8*,
 * rnnt_greedy_decode*) are ordered per device by the library itself — each waits (cudaStreamWaitEvent) for the one
 * launched before it, whatever stream that was on — because two grids that wait on flags of their own CTAs must never be
 * resident half each; every other kernel overlaps freely. */
int rnnt_check_device_status(void* stream);
/* Non-synchronising pair for pipelined callers (the SUT): enqueue a copy of the status word into *host_dst (pinned host
 * int32) on `stream`, look at it after the caller's own synchronisation point; non-zero = a watchdog fired and every
 * result produced since is invalid.  The word is per device and sticky: later kernels of this library on the device
 * bail out early until it is cleared (rnnt_clear_device_status, or rnnt_check_device_status). */
int rnnt_peek_device_status(int32_t* host_dst, void* stream);
int rnnt_clear_device_status(void* stream);

/* quant_tpp::quant_ps_epi8 + K zero-pad (mlperf_plugins/csrc/lstm_amx.cpp:35-50, tpps/quant_tpp.cpp:6-17)
 * in [rows][k_in] fp32 -> out [rows][k_out] int8, out[:,k_in:] = 0.  q = clamp(rne(x*scale),-128,127) */
int rnnt_quantize_pad_f32_i8(const float* in, int8_t* out, float scale, int64_t rows, int k_in, int k_out,
                             void* stream);

/* Same quantiser fed from the SUT's ragged batch (models/pytorch_sut.py:85-104 pad_sequence + batch pad, then
 * lstm_amx.cpp:35-50): packed [sum T_i][k_in] fp32, utterance n = rows [offsets[n], offsets[n]+lens[n]);
 * out {T,N,k_out} int8 time-major, zero beyond lens[n] and beyond k_in (lens[n] = 0 for batch-padding lanes). */
int rnnt_quantize_ragged_f32_i8(const float* packed, const int64_t* offsets, const int32_t* lens, int8_t* out, float scale, int T,
                                int N, int k_in, int k_out, void* stream);

/* i8o32b0 input projection over all timesteps (lstm_amx.cpp:96-113; tile_product_tpp.hpp:288-306):
 * C[M][N] int32 = A[M][K] int8 . W[N][K]^T int8, exact.  tcgen05 kind::i8, TMEM accumulators, TMA in/out.
 * Requirements: K % 16 == 0, N % 64 == 0, pointers 16-byte aligned. */
int rnnt_gemm_i8_i32(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, void* stream);
/* Same GEMM over a ragged time-major batch sorted by length: M = T * rows_per_step rows (any rows_per_step <= 4096);
 * batch tile b (rows [128 b, 128 b + 128) of a timestep) is live at timestep t if t < tile_steps_host[b] (HOST array,
 * ceil(rows_per_step / 128) entries).  A 128-row block of C whose rows are all dead is not computed and left untouched
 * (the recurrent kernel never reads dead rows); a block that straddles live and dead rows is computed whole.
 * tile_steps_host NULL = all rows. */
int rnnt_gemm_i8_i32_ragged(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, int rows_per_step,
                            const int32_t* tile_steps_host, void* stream);
/* Same product for N = 4096 = 4 gates x 1024 units, stored gate-blocked for rnnt_lstm_recurrent_i8_gate_blocked_ragged:
 * C[slice 32][gate 4][unit octet 4][row M][8] int32, element (row m, column gate*1024 + slice*32 + octet*8 + u) of the
 * row-major result.  Same values, same ragged skipping; C 32-byte aligned. */
int rnnt_gemm_i8_i32_gate_blocked_ragged(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, int rows_per_step,
                                         const int32_t* tile_steps_host, void* stream);
/* Same contract on CUDA cores (dp4a), M <= 65535.  TEST HELPER: lets the GPU tests cross-check the tcgen05
 * kernel at sizes the CPU oracle cannot reach.  Never called by the product path. */
int rnnt_gemm_i8_i32_simt(const int8_t* A, const int8_t* W, int32_t* C, int64_t M, int N, int K, void* stream);

/* Phase B of lstm_layer_amx_int8 (lstm_amx.cpp:122-152): for t in [0,T):
 *   gates = (float(h.W_hh^T + G[t]) + bias) * rb_scale ; lstm_postop_tpp::ref (tpps/lstm_postop_tpp.cpp:12-98)
 * as ONE persistent kernel.  G [T][N][4H] int32 (from rnnt_gemm_i8_i32), w_hh [4H][H], bias [4H] fp32,
 * hx [N][H] int8 and cx [N][H] binary16 updated IN PLACE, y [T][N][H] int8 (skip_quant_y=0) or bf16.
 * H must be 1024; N <= 4096 per call.  G must be 32-byte aligned (read with 256-bit loads), w_hh / hx / cx 16-byte, y 8-byte.
 * workspace: rnnt_lstm_recurrent_workspace_bytes(N) bytes, 256-byte aligned, contents don't care. */
int64_t rnnt_lstm_recurrent_workspace_bytes(int N);
int rnnt_lstm_recurrent_i8(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                           float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                           void* workspace, void* stream);
/* Same kernel for a ragged batch whose rows are sorted by length (PytorchSUT.issue_queries sorts descending,
 * models/pytorch_sut.py:58-65): tile_steps_host[b] (HOST array, ceil(N/128) entries, each in [0,T]) is the number of
 * timesteps batch tile b (rows [128b, 128b+128)) runs.  Steps t >= tile_steps[b] of that tile are skipped entirely:
 * y[t] is not written there and hx/cx hold the state after tile_steps[b] steps.  The reference runs every lane for all
 * T steps and discards the tail (stack_time.cpp:42-57 zeroes it, greedy_decode_update reads t < f_lens only), so
 * transcripts are unchanged as long as tile_steps[b] >= max length in the tile.  NULL = every tile runs T steps. */
int rnnt_lstm_recurrent_i8_ragged(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale, float in_scale,
                                  float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y, int T, int N, int H,
                                  const int32_t* tile_steps_host, void* workspace, void* stream);
/* Same kernel reading the gate-blocked projection written by rnnt_gemm_i8_i32_gate_blocked_ragged with M = T * N rows
 * (the 32 batch rows a warp handles are then 1 KB contiguous: 8 cache lines per load instead of 32). */
int rnnt_lstm_recurrent_i8_gate_blocked_ragged(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale,
                                               float in_scale, float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y,
                                               int T, int N, int H, const int32_t* tile_steps_host, void* workspace, void* stream);

/* rnnt_lstm_recurrent_i8 for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernel, arguments and results as rnnt_lstm_recurrent_i8, but no `workspace` argument: the scratch (device status
 * word, per-tile release counters, h ping-pong buffer, c scratch) is a file-scope device array inside the library —
 * 2,099,968 bytes that every GPU which loads the library holds until the process exits.  LIMIT: N <= 512 rows per call
 * (T unbounded); a larger N returns RNNT_ERR_INVALID with a message naming the limit.  The call performs no cudaMalloc /
 * cudaFree / allocating library routine.  The array is shared by all calls and keeps the last call's contents; what must
 * start at zero (status word, counters) is cleared on `stream` at the start of each call.  Calls that overlap on one
 * GPU (different streams, parallel graph branches) RACE on it: serialise them.  Device-side failures (watchdog) are
 * reported through status_out_host (may be NULL; page-locked host int32, written asynchronously on `stream`: 0 = ok)
 * — this variant does not touch the status word of rnnt_check_device_status / rnnt_peek_device_status. */
int rnnt_lstm_recurrent_i8_static_scratch(const int32_t* G, const int8_t* w_hh, const float* bias, float rb_scale,
                                          float in_scale, float out_scale, int skip_quant_y, int8_t* hx, uint16_t* cx, void* y,
                                          int T, int N, int H, int32_t* status_out_host, void* stream);

/* TUNING (process-wide, results do not depend on it): slice-CTAs per thread-block cluster of the recurrent kernel, 1, 2, 4
 * or 8 (default 1, or the environment variable RNNT_REC_CLUSTER); the h tile of a batch tile is TMA-multicast inside a
 * cluster.  Measured on B200: no gain over 1 (profiles/r01/rec_cluster_v5.txt). */
int rnnt_tune_recurrent_cluster(int ctas);
/* TUNING (process-wide, results do not depend on it): CTAs per cluster of the projection GEMM (1, 2 or 4; default 1, or the
 * environment variable RNNT_GEMM_CLUSTER).  The CTAs of a cluster compute consecutive m tiles of one n tile and share
 * its W operand by TMA multicast.  Measured on B200 (profiles/r02): no gain — the operand traffic does not bound the kernel. */
int rnnt_tune_gemm_cluster(int ctas);
/* TUNING (process-wide, results do not depend on it): run the projection GEMM on CTA pairs (tcgen05 cta_group::2, one
 * 256 x 256 tile per pair, each CTA staging half of the W tile) instead of single CTAs.  Default on (problems with at
 * least two m tiles and N % 256 == 0), or the environment variable RNNT_GEMM_PAIR=0 / 1. */
int rnnt_tune_gemm_pair(int on);

/* ---- audio front-end: the step before the hot path, datasets/parts/features.py:186-252.  fp32 throughout.
 * preemphasis (init.cpp: preemphasis; preemphasis.cpp:10-46): x [N][L], lens [N] (device int32) -> out [N][L + 2 pad]:
 * y[i] = x[i] - coeff * x[i-1] for i < len, reflect-padded by `pad` on both sides, zeros behind. */
int rnnt_preemphasis_f32(const float* x, const int32_t* lens, float coeff, int pad, float* out, int N, int L, void* stream);
/* power_spectrum (power_spectrum.cpp:10-33): x [N][T][F][2] (re, im) -> out [N][T][F] = re^2 + im^2, frames >= lens[n] zero. */
int rnnt_power_spectrum_f32(const float* x, const int32_t* lens, float* out, int N, int T, int F, void* stream);
/* frame_splicing (frame_splicing.cpp:12-45): x [N][C][T] -> out [N][C * factor][ceil(T / factor)],
 * out[n][k*C + c][j] = x[n][c][factor*j + k] if that frame is < lens[n], else 0. */
int rnnt_frame_splicing_f32(const float* x, const int32_t* lens, float* out, int N, int C, int T, int factor, void* stream);
/* i_layernorm_pad (normalization.cpp:10-64; i_layernorm_tpp.cpp:10-76): per (sample, feature) over the lens[n] valid frames:
 * var = E[x^2] - mean^2 + eps (times len/(len-1) if unbiased), y = (x - mean) / sqrt(var) * w + b; weight / bias [C][T] or
 * NULL (1 / 0).  out [n_pad][c_out][t_out]; everything the reference leaves unwritten (padding samples, features >= C,
 * frames >= len) is zero here. */
int rnnt_layernorm_pad_f32(const float* x, const int32_t* lens, const float* weight, const float* bias, float* out, int N, int C,
                           int T, int n_pad, int c_out, int t_out, float eps, int unbiased, void* stream);
/* torch.stft(n_fft 512, hop, window win_len centred in n_fft, center=False) + power_spectrum + dither^2 + mel filterbank +
 * log of features.py:201-227 in one kernel (the reference calls torch.stft, a library FFT, between its operators).
 * y [N][Lp] pre-emphasised, padded waveforms; frame_lens [N] (frames, device int32); window [win_len]; fb [n_mels][257];
 * out [N][n_mels][T] = log(fb . (|STFT|^2 masked past frame_lens + dither2) + log_bias).  n_fft must be 512, n_mels <= 96. */
int rnnt_logmel_f32(const float* y, const int32_t* frame_lens, const float* window, const float* fb, float* out, int N, int Lp,
                    int T, int n_fft, int hop, int win_len, int n_mels, float dither2, float log_bias, void* stream);

/* lstm_layer_amx_int8 (lstm_amx.cpp:59-158; schema init.cpp:129-131) = projection GEMM + recurrent kernel.
 * x [T][N][K] int8 (K % 16 == 0).  g_ws: T*N*4H int32 scratch for the projection. */
int rnnt_lstm_layer_i8(const int8_t* x, int8_t* hx, uint16_t* cx, const int8_t* w_ih, const int8_t* w_hh,
                       const float* bias, float rb_scale, float in_scale, float out_scale, int skip_quant_y, void* y,
                       int T, int N, int K, int H, int32_t* g_ws, void* workspace, void* stream);

/* stack_time (mlperf_plugins/csrc/stack_time.cpp:27-59; schema init.cpp:96-98), int8 only:
 * in {T,N,C} -> out {ceil(T/factor),N,C*factor}, zeros from lens[n] on. */
int rnnt_stack_time_i8(const int8_t* in, const int32_t* lens, int factor, int8_t* out, int T, int N, int C,
                       void* stream);

/* Transcription half of Joint fc1, hoisted out of the decode loop (amx_linear.cpp:137-150, i16o16b0):
 * y1[M][J] bf16 = bf16(f[M][C] . W_trans[J][C]^T), fp32 FMA sequential in k (the order the oracle defines). */
int rnnt_linear_bf16_bf16(const uint16_t* x, const uint16_t* w, uint16_t* y, int64_t M, int K, int O, void* stream);
/* Ragged variant (see rnnt_gemm_i8_i32_ragged; rows_per_step % 64 == 0): dead rows of y are left untouched. */
int rnnt_linear_bf16_bf16_ragged(const uint16_t* x, const uint16_t* w, uint16_t* y, int64_t M, int K, int O, int rows_per_step,
                                 const int32_t* tile_steps_host, void* stream);

/* The same product on the tensor cores (tcgen05.mma kind::f16, BF16 x BF16 -> FP32 in TMEM, bf16 RNE epilogue, TMA in/out):
 * the path the pipeline uses for the hoisted transcription half.  C[M][N] bf16 = bf16(A[M][K] . W[N][K]^T).  The order of
 * the fp32 accumulation inside the tensor core is not the oracle's (nor AMX's): results agree to <= 1 bf16 ulp (pinned
 * against AMX hardware output, tests/test_amx_golden.py).  K % 8 == 0, N % 64 == 0, 16-byte aligned pointers. */
int rnnt_gemm_bf16_bf16(const uint16_t* A, const uint16_t* W, uint16_t* C, int64_t M, int N, int K, void* stream);
/* Ragged variant (see rnnt_gemm_i8_i32_ragged): blocks of dead rows of C are left untouched. */
int rnnt_gemm_bf16_bf16_ragged(const uint16_t* A, const uint16_t* W, uint16_t* C, int64_t M, int N, int K, int rows_per_step,
                               const int32_t* tile_steps_host, void* stream);

/* lstm_postop (init.cpp:126-128; lstm_postop.cpp:14-72 -> tpps/lstm_postop_tpp.cpp:12-98) as a stand-alone operator:
 * it/ft/gt/ot fp32 [N][H] with a row pitch of `ld` floats (the four chunks of one [N][4H] tensor work in place);
 * ct [N][H] binary16 updated IN PLACE; h_q [N][H] int8 = quant(h, in_scale); y_q int8 = quant(h, out_scale) when
 * skip_quant_y == 0, else y_bf16 = bf16(h) (the other output pointer may be NULL).  fp16 gate math identical to the
 * fused recurrent kernel (Pade [5/6] + VRCPPH table).  H and ld even. */
int rnnt_lstm_postop_fp16(const float* it, const float* ft, const float* gt, const float* ot, int64_t ld, uint16_t* ct,
                          float in_scale, float out_scale, int skip_quant_y, uint16_t* y_bf16, int8_t* y_q, int8_t* h_q, int N,
                          int H, void* stream);

/* Per-iteration operators of the reference's decode loop, for callers that keep the Python loop of
 * models/decoder.py:188-211 (the fused rnnt_greedy_decode below is the fast path; these share its arithmetic).
 *   amx_linear_i16o32 (init.cpp:36-38; amx_linear.cpp:113-117): y[M][O] fp32 = x.W^T + bias (bias may be NULL). */
int rnnt_linear_bf16_f32(const uint16_t* x, const uint16_t* w, const float* bias, float* y, int64_t M, int K, int O,
                         void* stream);
/*   second half of amx_linear_bf16_accum_relu (init.cpp:33-35; amx_linear.cpp:152-168):
 *   y[M][O] bf16 = bf16(max((x1.W1^T + float(y1)) + bias, 0)), y1 = rnnt_linear_bf16_bf16(x0, W0). */
int rnnt_linear_bf16_accum_relu(const uint16_t* y1, const uint16_t* x1, const uint16_t* w1, const float* bias, uint16_t* y,
                                int64_t M, int K1, int O, void* stream);
/*   i8o32b32 dequant (amx_linear_i8o32, init.cpp:30-32; amx_linear.cpp:18-92): out = (float(acc) + bias[o]) * scale */
int rnnt_dequant_i32_f32(const int32_t* acc, const float* bias, float scale, float* out, int64_t M, int O, void* stream);
/*   lstm_layer_amx_bf16 at seq_len 1 (init.cpp:135-140; lstm_amx.cpp:187-264 + lstm_postop_tpp.cpp:100-163):
 *   x [N][K], h [N][P] bf16, c [N][P] fp32 -> NEW hy bf16, cy fp32.  bias = b_ih + b_hh.  gates_ws: 2*N*4P floats. */
int rnnt_lstm_cell_bf16(const uint16_t* x, const uint16_t* h, const float* c, const uint16_t* w_ih, const uint16_t* w_hh,
                        const float* bias, uint16_t* hy, float* cy, int N, int K, int P, float* gates_ws, void* stream);
/*   greedy_decode_update (init.cpp:141-143; greedy_decode_update.cpp:13-96).  State updated in place;
 *   *unfinished (device int32) receives the number of lanes with time_idx < f_lens (0 <=> the op's `true`). */
int rnnt_greedy_decode_update(const int64_t* symbols, int32_t* symbols_added, int32_t* res, int32_t* res_idx, const uint16_t* f,
                              const int32_t* f_lens, int32_t* time_idx, uint16_t* fi, int32_t* pre_g, uint16_t* pre_hg0,
                              uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, const uint16_t* hg0, const uint16_t* hg1,
                              const float* cg0, const float* cg1, int N, int L, int C, int P, int32_t* unfinished, void* stream);

/* Prediction + Joint + argmax + greedy_decode_update fused into one device-resident kernel
 * (models/decoder.py:171-212 loop; modeling_rnnt.py:183-205,259-289,335-365; greedy_decode_update.cpp:13-96).
 * Runs the whole while-loop on the device; per-utterance symbol counters, max_symbols_per_step = 30.
 * All model tensors bf16 row-major unless noted. */
typedef struct {
  const uint16_t* embed;      /* [28][P] bf16                                        */
  const uint16_t* w_ih[2];    /* [4P][P]                                             */
  const uint16_t* w_hh[2];    /* [4P][P]                                             */
  const float* bias[2];       /* [4P] = b_ih + b_hh  (modeling_rnnt.py:170-172)      */
  const uint16_t* w_pred;     /* [J][P]  joint.linear1_pred                          */
  const float* b1;            /* [J] = b_trans + b_pred (modeling_rnnt.py:225-227)   */
  const uint16_t* w2;         /* [32][J] rows 29..31 zero (modeling_rnnt.py:241-246) */
  const float* b2;            /* [32] entries 29..31 zero                            */
} rnnt_decode_weights;

int64_t rnnt_greedy_decode_workspace_bytes(int N);
/* y1 [T2][N][J] bf16 (from rnnt_gemm_bf16_bf16 or rnnt_linear_bf16_bf16), f_lens [N] STACKED lengths.
 * In/out state: res [N][L] int32, res_idx/symbols_added/time_idx/pre_g [N] int32,
 * pre_hg0/1 [N][P] bf16, pre_cg0/1 [N][P] fp32.  iters_out (device int32[1], may be NULL) receives
 * the number of batch-lockstep iterations the reference loop would have run.  N <= 2240 per call.
 * The three contractions of an iteration (prediction-LSTM gates, g.W_pred, fc2) run on the tensor cores (mma.sync bf16,
 * fp32 accumulation): their summation order is neither the CPU oracle's nor AMX's, so transcripts agree with either
 * except where the top-2 logits are a near-tie (tolerance + margin report).  Weights 16-byte aligned. */
int rnnt_greedy_decode(const uint16_t* y1, const int32_t* f_lens, int T2, int N, const rnnt_decode_weights* w_host,
                       int32_t* res, int L, int32_t* res_idx, int32_t* symbols_added, int32_t* time_idx, int32_t* pre_g,
                       uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1, int32_t* iters_out,
                       void* workspace, void* stream);
/* Same loop with the contractions on CUDA cores as fp32 FMA chains sequential in k — the accumulation order the CPU
 * oracle defines —, bit-exact against it: the cross-check of the state machine and of everything downstream of y1. */
int rnnt_greedy_decode_exact_order(const uint16_t* y1, const int32_t* f_lens, int T2, int N, const rnnt_decode_weights* w_host,
                                   int32_t* res, int L, int32_t* res_idx, int32_t* symbols_added, int32_t* time_idx,
                                   int32_t* pre_g, uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0, float* pre_cg1,
                                   int32_t* iters_out, void* workspace, void* stream);

#ifdef __cplusplus
}
#endif
#endif
