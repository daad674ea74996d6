// gemm_i8.cu — int8 x int8 -> int32 projection GEMM for sm_100a.
//
// Replaces phase A of lstm_layer_amx_int8 (mlperf_plugins/csrc/lstm_amx.cpp:96-113: T calls of the AMX
// tile kernel, epilogue i8o32b0 = raw int32 store, tpps/tile_product_tpp.hpp:288-306) with ONE launch over
// all timesteps: C[M = T*N][4H] = X[M][K] . W_ih[4H][K]^T.
//
// Design (persistent, warp-specialised, one CTA per SM):
//   warp 0   : TMA producer   — cp.async.bulk.tensor 128x128 B (A) and BNx128 B (W) boxes, SWIZZLE_128B,
//                               4-stage ring guarded by full/empty mbarriers
//   warp 1   : MMA issuer     — one elected lane issues tcgen05.mma.kind::i8 (M=128, N=BN, K=32) x4 per stage,
//                               accumulating in TMEM; 2 accumulator stages (2*BN <= 512 columns)
//   warp 2   : TMEM allocator
//   warps 4-11: epilogue      — tcgen05.ld 32 lanes x 32 columns -> swizzled smem staging -> TMA store of a
//                               32-row x 128-byte box per warp (fully coalesced int32 rows in HBM); two warps per TMEM
//                               lane quarter on alternate chunks (gemm_i8_epilogue.cuh)
// Clusters (template parameter CM = 1, 2, 4): the kernel is bound by the L2 -> SM operand traffic (ncu round 1: L2 71 %,
// tensor pipe 45 %, DRAM 45 %): a 128 x 256 tile reads (128 + 256) x K bytes.  CM CTAs of a thread-block cluster work on CM
// consecutive m tiles of the SAME n tile and share its W operand: each CTA fetches 1/CM of the W box and TMA-multicasts
// it into the same stage of every CTA of the cluster, so a CTA-tile costs (128 + 256 / CM) x K bytes of L2 reads.  A stage
// is free when the MMAs of ALL CM CTAs have consumed it (multicast tcgen05.commit onto every CTA's empty barrier).
// The int32 result is stored unscaled on purpose: the reference adds the recurrent int32 accumulator BEFORE
// converting to float (tile_product_tpp.hpp:336-342), and |acc| may exceed 2^24, so pre-scaling to float would
// break bit-exactness.  Bound: HBM write of C (16 B * 4H... = 4*N bytes per row) vs tensor pipe at large K.
#include "gemm_i8_epilogue.cuh"

namespace rnnt {


template <int BN, int CM>
__global__ void __launch_bounds__(GEMM_I8_THREADS, 1)
gemm_i8_i32_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                   const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles, int num_k_blocks,
                   int* __restrict__ status, const RowSkip skip, const int gate_blocked_and_debug, int32_t* __restrict__ c_direct,
                   const long long M, const int ldc) {
  using S = GemmSmem<BN>;
#ifdef RNNT_DEBUG_HOOKS
  const int dbg = gate_blocked_and_debug >> 8;        // profiling only: 1 no operand loads, 2 no MMAs, 4 no result stores, 8 every CTA stores all its tiles onto one tile of its own (L2-resident)
#else
  constexpr int dbg = 0;
#endif
  const int gate_blocked = gate_blocked_and_debug & 0xff;
  constexpr uint16_t MC_MASK = (uint16_t)((1u << CM) - 1u);
  const int rank = (CM > 1) ? (int)cluster_ctarank() : 0;
  const int cluster_id = blockIdx.x / CM, num_clusters = gridDim.x / CM;
  const int num_m_super = (num_m_tiles + CM - 1) / CM;
  const int num_super = num_m_super * num_n_tiles;
  // tile of this CTA inside super tile st: m tile (st / num_n) * CM + rank; it is `alive` unless past the matrix or a
  // dead (timestep, batch tile) of a ragged batch; a super tile is skipped by the whole cluster when no CTA is alive
  auto m_of = [&](int st, int r) { return (st / num_n_tiles) * CM + r; };
  auto alive_at = [&](int st, int r) { const int m = m_of(st, r); return m < num_m_tiles && !skip.dead((long long)m * GEMM_BM); };
  auto any_alive = [&](int st) { bool a = false; for (int r = 0; r < CM; ++r) a |= alive_at(st, r); return a; };
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + S::STAGES * S::A_BYTES;
  const uint32_t smem_c = smem_base + S::STAGES * S::STAGE_BYTES;
  const uint32_t bar_base = smem_base + S::BAR_OFFSET;
  const uint32_t full_bar = bar_base;                 // [STAGES]
  const uint32_t empty_bar = bar_base + 8 * S::STAGES;
  const uint32_t tfull_bar = bar_base + 16 * S::STAGES;       // [2]
  const uint32_t tempty_bar = tfull_bar + 16;                   // [2]
  const uint32_t tmem_slot = tempty_bar + 16;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S::STAGES; ++i) { mbar_init(full_bar + 8 * i, 1); mbar_init(empty_bar + 8 * i, CM); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull_bar + 8 * i, 1); mbar_init(tempty_bar + 8 * i, GEMM_I8_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<2 * BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  if (CM > 1) cluster_sync_all();     // peers' barriers are initialised before any multicast / remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *(volatile uint32_t*)(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; bool ok = true;
      for (int st = cluster_id; st < num_super && ok; st += num_clusters) {
        if (!any_alive(st)) continue;
        const bool alive = alive_at(st, rank);
        const int m_blk = m_of(st, rank), n_blk = st % num_n_tiles;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          // the stage is free in EVERY CTA of the cluster once this barrier flips (multicast commits)
          if (!mbar_wait(empty_bar + 8 * stage, phase ^ 1, status)) { ok = false; break; }
          if (dbg & 1) { mbar_arrive(full_bar + 8 * stage); if (++stage == S::STAGES) { stage = 0; phase ^= 1; } continue; }
          mbar_expect_tx(full_bar + 8 * stage, (alive ? S::A_BYTES : 0) + S::B_BYTES);
          if (alive) tma_load_2d(smem_a + stage * S::A_BYTES, &tmap_a, full_bar + 8 * stage, kb * GEMM_BK, m_blk * GEMM_BM);
          if (CM == 1) {
            tma_load_2d(smem_b + stage * S::B_BYTES, &tmap_b, full_bar + 8 * stage, kb * GEMM_BK, n_blk * BN);
          } else {        // this CTA's 1/CM of the W box, into the same place of every CTA's stage
            tma_load_2d_mc(smem_b + stage * S::B_BYTES + rank * (BN / CM) * GEMM_BK, &tmap_b, full_bar + 8 * stage, kb * GEMM_BK,
                           n_blk * BN + rank * (BN / CM), MC_MASK);
          }
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_i8(GEMM_BM, BN);
      int stage = 0; uint32_t phase = 0; int acc = 0; uint32_t acc_phase = 0; bool ok = true;
      for (int st = cluster_id; st < num_super && ok; st += num_clusters) {
        if (!any_alive(st)) continue;
        const bool alive = alive_at(st, rank);
        if (alive) {
          if (!mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1, status)) break;
          tc_fence_after();
        }
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          if (!mbar_wait(full_bar + 8 * stage, phase, status)) { ok = false; break; }
          tc_fence_after();
          if (alive && !(dbg & 2)) {
            const uint64_t adesc = umma_desc_sw128(smem_a + stage * S::A_BYTES);
            const uint64_t bdesc = umma_desc_sw128(smem_b + stage * S::B_BYTES);
#pragma unroll
            for (int k = 0; k < GEMM_BK / 32; ++k)
              umma_i8(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          }
          // frees the stage (in every CTA of the cluster) when the MMAs issued so far have retired; a CTA without a live
          // tile still takes part: its peers multicast into its stage too
          if (CM == 1) umma_commit(empty_bar + 8 * stage);
          else umma_commit_mc(empty_bar + 8 * stage, MC_MASK);
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
        if (!ok) break;
        if (alive) {
          umma_commit(tfull_bar + 8 * acc);       // accumulator complete -> epilogue
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (TMEM -> smem -> TMA store)
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const int half = (warp - 4) >> 2;             // the two warps of a quarter take alternate 32-column chunks
    const uint32_t cbuf0 = smem_c + (uint32_t)((warp - 4) * 2) * S::C_WARP_BYTES;
    int acc = 0; uint32_t acc_phase = 0; int buf = 0;
    for (int st = cluster_id; st < num_super; st += num_clusters) {
      if (!alive_at(st, rank)) continue;
      const int m_blk = m_of(st, rank), n_blk = st % num_n_tiles;
      if (!mbar_wait(tfull_bar + 8 * acc, acc_phase, status)) break;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if (!c_direct) {
        gemm_i8_store_chunks(t_row, cbuf0, buf, half, GEMM_I8_EPI_WARPS / 4, BN / 32, &tmap_c, gate_blocked,
                             (dbg & 8) ? (int)blockIdx.x * GEMM_BM + q * 32 : m_blk * GEMM_BM + q * 32, (dbg & 8) ? 0 : n_blk * BN, lane, (dbg & 4) != 0);
      } else {
#pragma unroll 1
        for (int c = half; c < BN / 32; c += GEMM_I8_EPI_WARPS / 4) {
          uint32_t v[32];
          RNNT_TMEM_LD_X16(t_row + c * 32, v);
          RNNT_TMEM_LD_X16(t_row + c * 32 + 16, v + 16);
          tmem_ld_wait();
          if (dbg & 4) continue;
          // straight from registers to global memory: a thread owns 32 consecutive columns of one row = one full 128-byte
          // line (row-major) or four 32-byte sectors, each part of a 1 KB run written by the warp (gate-blocked).
          // Measured slower than the staged TMA stores.
          const long long row = (long long)m_blk * GEMM_BM + q * 32 + lane;
          if (row < M) {
            const int col0 = n_blk * BN + c * 32;
            if (!gate_blocked) {
              int4* dst = reinterpret_cast<int4*>(c_direct + row * ldc + col0);
#pragma unroll
              for (int j = 0; j < 8; ++j) dst[j] = make_int4((int)v[4 * j], (int)v[4 * j + 1], (int)v[4 * j + 2], (int)v[4 * j + 3]);
            } else {
              const int gate = col0 >> 10, slice = (col0 & 1023) >> 5;
#pragma unroll
              for (int o = 0; o < 4; ++o) {
                int4* dst = reinterpret_cast<int4*>(c_direct + ((((long long)slice * 4 + gate) * 4 + o) * M + row) * 8);
                dst[0] = make_int4((int)v[8 * o], (int)v[8 * o + 1], (int)v[8 * o + 2], (int)v[8 * o + 3]);
                dst[1] = make_int4((int)v[8 * o + 4], (int)v[8 * o + 5], (int)v[8 * o + 6], (int)v[8 * o + 7]);
              }
            }
          }
        }
      }
      // all TMEM reads of this accumulator are complete (wait::ld above): hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (CM > 1) cluster_sync_all();     // no CTA leaves while peers may still multicast into it / arrive on its barriers
  if (warp == 2) { tc_fence_after(); tmem_dealloc<2 * BN>(tmem_base); }
}

// explicit instantiations used by api.cu
#define RNNT_GEMM_INST(BN, CM)                                                                                              \
  template __global__ void gemm_i8_i32_kernel<BN, CM>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, \
                                                      const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__,  \
                                                      const RowSkip, const int, int32_t* __restrict__, const long long, const int);
RNNT_GEMM_INST(256, 1) RNNT_GEMM_INST(256, 2) RNNT_GEMM_INST(256, 4) RNNT_GEMM_INST(128, 1) RNNT_GEMM_INST(64, 1)

// ---------------------------------------------------------------------------------------------
// Plain CUDA-core int8 GEMM.  NOT a fallback: used by tests (device-side cross-check of the tcgen05 kernel at
// sizes the CPU oracle cannot reach) and exposed as rnnt_gemm_i8_i32_simt for that purpose only.
__global__ void gemm_i8_i32_simt_kernel(const int8_t* __restrict__ A, const int8_t* __restrict__ W, int32_t* __restrict__ C,
                                        long long M, int N, int K) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const long long m = blockIdx.y;
  if (n >= N || m >= M) return;
  const int* a = reinterpret_cast<const int*>(A + m * K);
  const int* w = reinterpret_cast<const int*>(W + (long long)n * K);
  int acc = 0;
  for (int k = 0; k < K / 4; ++k) acc = __dp4a(a[k], w[k], acc);
  C[m * N + n] = acc;
}

}  // namespace rnnt
