// kernels.cuh — tile constants, shared-memory plans, parameter blocks and kernel declarations shared between
// the kernel translation units and api.cu.
#pragma once
#include "common.cuh"

// Profiling hooks (cycle traces, switch-off flags) exist only in the debug build of the library (`make debug`,
// -DRNNT_DEBUG_HOOKS, include/rnnt_b200_debug.h); in the product build these fold to constants and the branches vanish.
#ifdef RNNT_DEBUG_HOOKS
#define RNNT_DBG(p, bits) ((p).debug & (bits))
#define RNNT_TRACE_PTR(p) ((p).trace)
#else
#define RNNT_DBG(p, bits) 0
#define RNNT_TRACE_PTR(p) ((long long*)nullptr)
#endif

namespace rnnt {

// Ragged batches sorted by length: rows are (t, n) with rows_per_t = N rows per timestep; batch tile b = rows
// [128 b, 128 b + 128) of every timestep is live for t < steps[b].  rows_per_t == 0 switches the test off.
// dead(m0, rows): are ALL of the rows [m0, m0 + rows) dead?  Exact for any N: a block of rows that straddles batch
// tiles or timesteps is walked segment by segment (all rows of a segment share (t, b)); rows past the matrix are dead.
struct RowSkip {
  int rows_per_t;
  int steps[32];
  __device__ __forceinline__ bool dead(long long m0, int rows = 128) const {
    if (rows_per_t == 0) return false;
    const long long end = m0 + rows;
    long long m = m0;
    while (m < end) {
      const int t = (int)(m / rows_per_t), n = (int)(m % rows_per_t), b = n >> 7;
      if (t < steps[b]) return false;
      m += min((b + 1) << 7, rows_per_t) - n;
    }
    return true;
  }
};

// ------------------------------------------------------------------ gemm_i8.cu
constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 128;   // bytes of K per stage = one 128-byte swizzle row
constexpr int GEMM_STAGES = 4;      // gemm_bf16.cu
constexpr int GEMM_THREADS = 256;   // gemm_bf16.cu: 4 control + 4 epilogue warps
constexpr int GEMM_I8_EPI_WARPS = 8;                            // two per TMEM lane quarter (gemm_i8_epilogue.cuh)
constexpr int GEMM_I8_THREADS = 32 * (4 + GEMM_I8_EPI_WARPS);   // 384

template <int BN>
struct GemmSmem {
  static constexpr int STAGES = BN == 256 ? 3 : 4;
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK;           // 16 KB
  static constexpr int B_BYTES = BN * GEMM_BK;                // 32 KB at BN=256
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int C_WARP_BYTES = 32 * 128;               // 32 rows x 32 int32
  static constexpr int C_BYTES = GEMM_I8_EPI_WARPS * 2 /*buffers*/ * C_WARP_BYTES;  // 64 KB
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + C_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;       // barriers + alignment slack
};

template <int BN, int CM>
__global__ void gemm_i8_i32_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                                   const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles,
                                   int num_k_blocks, int* __restrict__ status, const RowSkip skip, const int gate_blocked,
                                   int32_t* __restrict__ c_direct, const long long M, const int ldc);
// gemm_i8_pair.cu: one 256 x 256 tile per CTA pair (cta_group::2); each CTA stages 128 rows of A and 128 rows of W
struct GemmPairSmem {
  static constexpr int STAGES = 5;
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK;           // 16 KB
  static constexpr int B_BYTES = 128 * GEMM_BK;               // 16 KB: this CTA's half of the 256-row W tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int C_WARP_BYTES = 32 * 128;
  static constexpr int C_BYTES = GEMM_I8_EPI_WARPS * 2 * C_WARP_BYTES;   // 64 KB
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + C_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;       // 225.25 KB
};
__global__ void gemm_i8_i32_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                                        const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles,
                                        int num_k_blocks, int* __restrict__ status, const RowSkip skip, const int gate_blocked);
__global__ void gemm_i8_i32_simt_kernel(const int8_t* __restrict__ A, const int8_t* __restrict__ W, int32_t* __restrict__ C,
                                        long long M, int N, int K);

// ------------------------------------------------------------------ gemm_bf16.cu
template <int BN>
struct GemmBf16Smem {
  static constexpr int BK_ELEMS = 64;                         // bf16 elements of K per stage = one 128-byte swizzle row
  static constexpr int A_BYTES = GEMM_BM * 128;               // 16 KB
  static constexpr int B_BYTES = BN * 128;                    // 32 KB at BN=256
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int C_WARP_BYTES = 32 * 64;                // 32 rows x 32 bf16
  static constexpr int C_BYTES = 4 /*warps*/ * 2 /*buffers*/ * C_WARP_BYTES;  // 16 KB
  static constexpr int BAR_OFFSET = GEMM_STAGES * STAGE_BYTES + C_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;       // barriers + alignment slack
};
template <int BN>
__global__ void gemm_bf16_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                                      const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles,
                                      int num_k_blocks, int* __restrict__ status, const RowSkip skip);

// ------------------------------------------------------------------ lstm_recurrent.cu
constexpr int REC_H = 1024;
constexpr int REC_BM = 128;          // batch rows per tile (UMMA M)
constexpr int REC_HS = 32;           // hidden units per slice CTA (all four gates: UMMA N = 128)
constexpr int REC_KC = 8;            // K chunks of 128 bytes per tile
constexpr int REC_TAB_N = 0x44FA;    // entries of the fp16 activation table: |x| bit patterns 0 .. 0x44F9 (4.97 = 0x44F8)
constexpr int REC_CTRL_WARPS = 4;    // 0: TMA producer, 1: MMA issuer, 2: TMEM alloc, 3: publisher
constexpr int REC_EPI_WARPS = 16;
constexpr int REC_STORE_WARPS = 4;   // epilogue warps 0..3 also write the staged h / y rows to global memory
constexpr int REC_THREADS = 32 * (REC_CTRL_WARPS + REC_EPI_WARPS);   // 640
constexpr int REC_FLAG_STRIDE = 32;  // ints between per-tile counters (128 B)
constexpr int REC_MAX_TILES = 32;    // batch tiles per launch (N <= 4096)
constexpr int REC_GRID_Y = 4;        // batch-tile groups: 32 slices x 4 = 128 CTAs

// HS = hidden units per slice CTA.  32 (UMMA N = 128; 32 slices x 4 tile groups) is the throughput configuration.  16 (UMMA
// N = 64; 64 slices x 2 tile groups) serves batches of <= 4 tiles, where every tile group runs ONE dependency chain per
// timestep: a thread's gate math covers 4 units instead of 8, and half the W slice leaves room for 32 KB ring slots, i.e.
// half as many wait / commit hand-shakes of the MMA warp per item (profiles/r02/rec_trace_chain_mode.txt).
template <int HS>
struct RecSmemT {
  static constexpr int CPS = HS == 32 ? 1 : 2;            // 16 KB K-chunks per ring slot: one wait / commit hand-shake of the MMA warp per slot
  static constexpr int NS = 3;                            // slots of the h operand ring (16 | 32 KB each)
  static constexpr int UMMA_N = 4 * HS;                   // 128 | 64
  static constexpr int UPT = HS / 4;                      // hidden units per epilogue thread: 8 | 4
  static constexpr int W_CHUNK = UMMA_N * 128;            // bytes per K-chunk of the weight slice (16 | 8 KB)
  static constexpr int W_BYTES = REC_KC * W_CHUNK;        // 128 | 64 KB
  static constexpr int H_CHUNK = REC_BM * 128;            // 16 KB
  static constexpr int SLOT = CPS * H_CHUNK;              // 16 | 32 KB
  static constexpr int H_BYTES = NS * SLOT;               // 48 | 96 KB
  static constexpr int OUT_OFFSET = W_BYTES + H_BYTES;    // staging of h (int8) and y (int8 | bf16) rows, one item deep;
  static constexpr int OUT_H = 8 * REC_BM * 4;            // [unit quad][row] 4 B  = 4 KB     (the rcp LUT lives here during set-up)
  static constexpr int OUT_Y = 8 * REC_BM * 8;            // [unit quad][row] 8 B  = 8 KB
  static constexpr int OUT_STAGE = OUT_H + OUT_Y;
  static constexpr int LUT_OFFSET = OUT_OFFSET;
  static constexpr int BIAS_OFFSET = OUT_OFFSET + OUT_STAGE;         // [4 gates][HS units] fp32
  static constexpr int TAB_OFFSET = BIAS_OFFSET + 512;               // whole-function table of the Pade tanh kernel
  static constexpr int TAB_BYTES = (REC_TAB_N * 2 + 15) / 16 * 16;   // 35,328
  static constexpr int BAR_OFFSET = TAB_OFFSET + TAB_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 512;          // barriers/counters + two int[32] tables (the window itself is 1024-byte aligned)
};
using RecSmem = RecSmemT<REC_HS>;
static_assert(RecSmemT<16>::TOTAL <= 232448, "recurrent kernel shared memory (HS = 16)");
static_assert(RecSmem::TOTAL <= 232448, "recurrent kernel shared memory");

struct RecParams {
  const int32_t* g;        // int32 input projection, read straight from global (one 32-byte sector per thread and gate):
  int g_blocked;           //   0: [T][N][4H] row-major   1: [32 slices][4 gates][4 unit octets][T*N rows][8] (coalesced per warp)
  const float* bias;       // [4H]
  int8_t* h_buf;           // [2][tiles][8 chunks][128 rows][128 B] ping-pong (workspace), K-chunk-major
  int8_t* hx;              // [N][H] caller's state: h_0 on entry (read by TMA), receives h after the last step of each tile
  uint16_t* cx;            // [N][H] binary16: c_0 on entry, final c on exit
  uint16_t* c_ws;          // [tiles][32 slices][4][128 rows][8] binary16: c between steps, CTA-private blocked layout
  void* y;                 // [T][N][H] int8 or bf16
  int* flags;              // [num_btiles][REC_FLAG_STRIDE] zeroed
  int* status;
  float rb_scale, in_scale, out_scale;
  int T, N, num_btiles, skip_quant_y;
  int tile_steps[REC_MAX_TILES];   // timesteps to run for each batch tile (<= T); ragged batches sorted by length
  int group_count[REC_GRID_Y];     // batch tiles of each of the grid's tile groups (blockIdx.y) ...
  int group_tile[REC_GRID_Y][REC_MAX_TILES];   // ... and their ids
  int debug;               // profiling only (rnnt_debug_set_recurrent_flags): 1 skip MMAs, 2 skip gate math, 4 skip h loads, 8 skip y/c stores,
                           // 16 relaxed publish, 32 no proxy fence, 64 no G, 128 no c prefetch, 256 ignore the h_t flags
  long long* trace;        // optional [T][16] clock64 stamps of CTA (0,0) (profiles/trace_recurrent.py; NULL otherwise)
};

template <int CS, int HS = REC_HS>
__global__ void lstm_recurrent_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_h0,
                                      const __grid_constant__ CUtensorMap tmap_h, const RecParams p);

// ------------------------------------------------------------------ elementwise.cu
constexpr int LB_TM = 64, LB_TO = 64, LB_TK = 32;
constexpr int LBB_T = 128, LBB_TK = 32;     // large-M variant: 128 x 128 tile, 8 x 8 outputs per thread
__global__ void quantize_pad_kernel(const float* __restrict__ in, int8_t* __restrict__ out, float scale, long long rows,
                                    int k_in, int k_out);
__global__ void quantize_ragged_kernel(const float* __restrict__ packed, const long long* __restrict__ offsets,
                                       const int32_t* __restrict__ lens, int8_t* __restrict__ out, float scale, int T, int N,
                                       int k_in, int k_out);
__global__ void stack_time_kernel(const int8_t* __restrict__ in, const int32_t* __restrict__ lens, int factor,
                                  int8_t* __restrict__ out, int T, int N, int C);
enum : int { LIN_BF16 = 0, LIN_F32_BIAS = 1, LIN_F32 = 2, LIN_ACCUM_RELU = 3 };
template <int MODE>
__global__ void linear_bf16_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w,
                                   const float* __restrict__ bias, const uint16_t* __restrict__ prev, void* __restrict__ y,
                                   long long M, int K, int O, const RowSkip skip);
template <int MODE>
__global__ void linear_bf16_big_kernel(const uint16_t* __restrict__ x, const uint16_t* __restrict__ w,
                                       const float* __restrict__ bias, const uint16_t* __restrict__ prev, void* __restrict__ y,
                                       long long M, int K, int O, const RowSkip skip);
__global__ void dequant_i32_f32_kernel(const int32_t* __restrict__ acc, const float* __restrict__ bias, float scale,
                                       float* __restrict__ out, long long total, int O);
__global__ void lstm_postop_f32_kernel(const float* __restrict__ gi, const float* __restrict__ gh, const float* __restrict__ bias,
                                       const float* __restrict__ c_in, uint16_t* __restrict__ hy, float* __restrict__ cy,
                                       int N, int P);
__global__ void lstm_postop_fp16_kernel(const float* __restrict__ it, const float* __restrict__ ft, const float* __restrict__ gt,
                                        const float* __restrict__ ot, long long ld, uint16_t* __restrict__ ct, float in_scale,
                                        float out_scale, int skip_quant_y, uint16_t* __restrict__ y_bf16,
                                        int8_t* __restrict__ y_q, int8_t* __restrict__ h_q, int N, int H);
__global__ void rec_copy_h1_kernel(const int8_t* __restrict__ h_buf, int8_t* __restrict__ hx, int num_btiles, int N,
                                   unsigned mask);
__global__ void greedy_update_kernel(const long long* __restrict__ symbols, int32_t* symbols_added, int32_t* res, int32_t* res_idx,
                                     const uint16_t* __restrict__ f, const int32_t* __restrict__ f_lens, int32_t* time_idx,
                                     uint16_t* fi, int32_t* pre_g, uint16_t* pre_hg0, uint16_t* pre_hg1, float* pre_cg0,
                                     float* pre_cg1, const uint16_t* __restrict__ hg0, const uint16_t* __restrict__ hg1,
                                     const float* __restrict__ cg0, const float* __restrict__ cg1, int N, int L, int C, int P,
                                     int* unfinished);

// ------------------------------------------------------------------ greedy_decode.cu
constexpr int DEC_P = 320, DEC_J = 512, DEC_V = 32;
constexpr int DEC_S = 20;             // CTAs per lane group
constexpr int DEC_U = DEC_P / DEC_S;  // 16 hidden units per CTA
constexpr int DEC_COLS = 4 * DEC_U;   // 64 gate columns per CTA and matrix
constexpr int DEC_KSTR = DEC_P + 4;   // padded bf16 row stride (648 B: conflict-free 8-byte reads)
constexpr int DEC_THREADS = 256;
constexpr int DEC_WARPS = DEC_THREADS / 32;
constexpr int DEC_CHUNK = 12;         // active lanes per LSTM pass (the fc1 pass takes 2 * DEC_CHUNK)
constexpr int DEC_JC = 26;            // fc1 columns per CTA (19*26 + 18 = 512)
constexpr int DEC_MAXLG = 320;        // max lanes per group (N <= 7*320)
constexpr int DEC_BLANK = 28, DEC_MAXSYM = 30;

struct DecSmem {
  static constexpr int W_BYTES = 2 * 2 * DEC_COLS * DEC_KSTR * 2;          // [layer][ih|hh][col][k] bf16 = 165,888
  static constexpr int ACT_BYTES = DEC_CHUNK * 2 * DEC_P * 4;              // [lane][x|h] fp32 = 30,720 (joint: per-warp y)
  static constexpr int PART_BYTES = 2 * DEC_CHUNK * DEC_COLS * 4;          // [ih|hh][lane][col] fp32 = 6,144
  static constexpr int PAD_BYTES = 1536;                                   // rounds the joint-phase overlay up to 38,400
  static constexpr int WP_BYTES = DEC_JC * DEC_KSTR * 2;                   // this CTA's W_pred rows, bf16 = 16,848
  // joint phase overlays ACT|PART|PAD: fc2 rows 0..28 (bf16, row stride 1040 B) + one bf16 y vector per warp
  static constexpr int W2_ROWS = 29, W2_STRIDE = DEC_J * 2 + 16;
  static constexpr int JY_OFFSET = W2_ROWS * W2_STRIDE;                    // 30,160
  static_assert(JY_OFFSET + DEC_WARPS * DEC_J * 2 <= ACT_BYTES + PART_BYTES + PAD_BYTES, "joint overlay does not fit");
  static constexpr int TOTAL = W_BYTES + ACT_BYTES + PART_BYTES + PAD_BYTES + WP_BYTES;  // 221,136 (+ ~9 KB static)
};

struct DecParams {
  // model
  const uint16_t* embed; const uint16_t* w_ih[2]; const uint16_t* w_hh[2]; const float* bias[2];
  const uint16_t* w_pred; const float* b1; const uint16_t* w2; const float* b2;
  // inputs
  const uint16_t* y1; const int32_t* f_lens; int T2, N;
  // in/out state
  int32_t* res; int L; int32_t* res_idx; int32_t* symbols_added; int32_t* time_idx; int32_t* pre_g;
  uint16_t* hg[2][2];   // [buffer][layer] : buffer 0 = caller's pre_hg arrays, buffer 1 = workspace
  float* cg[2][2];
  // workspace
  float* z;             // [N][J] fp32 : g.W_pred of the current candidate
  int32_t* lane_state;  // [2][N][4]: flags (1 cur, 2 need_lstm, 4 finished, 8 fixed), cur_t, pre_g, fix_iter; double-buffered by round parity
  int32_t* group_cnt;   // [NG][32] barrier arrival counters
  int32_t* global_sync; // [0] groups done, [1] max iterations
  int32_t* iters_out; int* status;
  int NG, LG;           // groups, lanes per group
  int max_run;          // decisions a lane may take per round without a group barrier (run-ahead through blank frames)
  long long* trace;     // optional [iters][8] clock64 stamps of CTA 0 (profiles/trace_decode.py; NULL otherwise)
  int trace_cap;
  int trace_block;      // the CTA whose thread 0 writes the trace (debug build)
  int skew;             // debug build: CTA 1 of every lane group idles this many cycles before each scan (tests/test_gpu_decode_skew.py)
};

// tensor-core variant: bf16 weights / activations as XOR-swizzled 16-byte chunks (conflict-free ldmatrix), no row padding
constexpr int DEC_TC_LANES = 32;      // active lanes per LSTM / fc1 pass (two m16 tiles)
struct DecSmemTC {
  static constexpr int W_BYTES = 2 * 2 * DEC_COLS * DEC_P * 2;             // [layer][ih|hh][col][k] bf16 = 163,840
  static constexpr int ACT_BYTES = DEC_TC_LANES * 2 * DEC_P * 2;           // [lane][x|h] bf16 = 40,960 (then the fp32 gate sums)
  static constexpr int WP_BYTES = DEC_JC * DEC_P * 2;                      // this CTA's W_pred rows = 16,640
  // joint phase overlays act: y of the CTA's <= 16 lanes (bf16 rows of 512, pitch 1040 B) + fp32 logits [16][32]
  static constexpr int JY_PITCH = DEC_J * 2 + 16;
  static constexpr int JY_BYTES = 16 * JY_PITCH;                           // 16,640
  static_assert(JY_BYTES + 2 * 16 * DEC_V * 4 <= ACT_BYTES, "joint overlay does not fit");
  static_assert(DEC_MAXLG <= 16 * DEC_S, "a CTA's lanes must fit one m16 tile");
  static_assert(2 * DEC_TC_LANES * DEC_COLS * 4 <= ACT_BYTES, "gate sums do not fit");
  static constexpr int TOTAL = W_BYTES + ACT_BYTES + WP_BYTES;             // 221,440 (+ ~9 KB static)
};

template <bool TC>
__global__ void greedy_decode_kernel(const DecParams p);

}  // namespace rnnt
