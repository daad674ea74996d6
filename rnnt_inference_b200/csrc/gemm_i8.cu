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
//   warps 4-7: epilogue       — tcgen05.ld 32 lanes x 32 columns -> swizzled smem staging -> TMA store of a
//                               32-row x 128-byte box per warp (fully coalesced int32 rows in HBM)
// The int32 result is stored unscaled on purpose: the reference adds the recurrent int32 accumulator BEFORE
// converting to float (tile_product_tpp.hpp:336-342), and |acc| may exceed 2^24, so pre-scaling to float would
// break bit-exactness.  Bound: HBM write of C (16 B * 4H... = 4*N bytes per row) vs tensor pipe at large K.
#include "kernels.cuh"

namespace rnnt {


template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_i8_i32_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                   const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles, int num_k_blocks,
                   int* __restrict__ status, const RowSkip skip, const int gate_blocked) {
  using S = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + GEMM_STAGES * S::A_BYTES;
  const uint32_t smem_c = smem_base + GEMM_STAGES * S::STAGE_BYTES;
  const uint32_t bar_base = smem_base + S::BAR_OFFSET;
  const uint32_t full_bar = bar_base;                 // [STAGES]
  const uint32_t empty_bar = bar_base + 8 * GEMM_STAGES;
  const uint32_t tfull_bar = bar_base + 16 * GEMM_STAGES;       // [2]
  const uint32_t tempty_bar = tfull_bar + 16;                   // [2]
  const uint32_t tmem_slot = tempty_bar + 16;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = num_m_tiles * num_n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < GEMM_STAGES; ++i) { mbar_init(full_bar + 8 * i, 1); mbar_init(empty_bar + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull_bar + 8 * i, 1); mbar_init(tempty_bar + 8 * i, 4); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<2 * BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *(volatile uint32_t*)(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; bool ok = true;
      for (int tile = blockIdx.x; tile < num_tiles && ok; tile += gridDim.x) {
        const int m_blk = tile / num_n_tiles, n_blk = tile % num_n_tiles;
        if (skip.dead((long long)m_blk * GEMM_BM)) continue;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          if (!mbar_wait(empty_bar + 8 * stage, phase ^ 1, status)) { ok = false; break; }
          mbar_expect_tx(full_bar + 8 * stage, S::STAGE_BYTES);
          tma_load_2d(smem_a + stage * S::A_BYTES, &tmap_a, full_bar + 8 * stage, kb * GEMM_BK, m_blk * GEMM_BM);
          tma_load_2d(smem_b + stage * S::B_BYTES, &tmap_b, full_bar + 8 * stage, kb * GEMM_BK, n_blk * BN);
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_i8(GEMM_BM, BN);
      int stage = 0; uint32_t phase = 0; int acc = 0; uint32_t acc_phase = 0; bool ok = true;
      for (int tile = blockIdx.x; tile < num_tiles && ok; tile += gridDim.x) {
        if (skip.dead((long long)(tile / num_n_tiles) * GEMM_BM)) continue;
        if (!mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1, status)) break;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          if (!mbar_wait(full_bar + 8 * stage, phase, status)) { ok = false; break; }
          tc_fence_after();
          const uint64_t adesc = umma_desc_sw128(smem_a + stage * S::A_BYTES);
          const uint64_t bdesc = umma_desc_sw128(smem_b + stage * S::B_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 32; ++k)
            umma_i8(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          umma_commit(empty_bar + 8 * stage);   // frees the smem stage when these MMAs retire
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
        if (!ok) break;
        umma_commit(tfull_bar + 8 * acc);       // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (TMEM -> smem -> TMA store)
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const uint32_t cbuf0 = smem_c + (uint32_t)(q * 2) * S::C_WARP_BYTES;
    int acc = 0; uint32_t acc_phase = 0; int buf = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / num_n_tiles, n_blk = tile % num_n_tiles;
      if (skip.dead((long long)m_blk * GEMM_BM)) continue;
      if (!mbar_wait(tfull_bar + 8 * acc, acc_phase, status)) break;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        RNNT_TMEM_LD_X16(t_row + c * 32, v);
        RNNT_TMEM_LD_X16(t_row + c * 32 + 16, v + 16);
        tmem_ld_wait();
        // staging buffer `buf` was last read by the TMA store issued two chunks ago
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        if (!gate_blocked) {
          const uint32_t dst = cbuf0 + (uint32_t)buf * S::C_WARP_BYTES + (uint32_t)lane * 128u;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t a = dst + (uint32_t)((j ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v[4 * j]), "r"(v[4 * j + 1]),
                         "r"(v[4 * j + 2]), "r"(v[4 * j + 3]) : "memory");
          }
        } else {
          // gate-blocked result (see api.cu): staging = [unit octet 4][row 32][8 units] with the 32-byte TMA swizzle
          // (16-byte half ^= bit 2 of the row), which spreads the eight lanes of a store phase over all 32 banks
          const uint32_t dst = cbuf0 + (uint32_t)buf * S::C_WARP_BYTES + (uint32_t)lane * 32u;
          const uint32_t hsw = (uint32_t)((lane >> 2) & 1);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t a = dst + (uint32_t)((j >> 1) * 1024) + ((((uint32_t)j & 1u) ^ hsw) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v[4 * j]), "r"(v[4 * j + 1]),
                         "r"(v[4 * j + 2]), "r"(v[4 * j + 3]) : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          const int col0 = n_blk * BN + c * 32, row0 = m_blk * GEMM_BM + q * 32;
          if (!gate_blocked) tma_store_2d(&tmap_c, cbuf0 + (uint32_t)buf * S::C_WARP_BYTES, col0, row0);
          else tma_store_5d(&tmap_c, cbuf0 + (uint32_t)buf * S::C_WARP_BYTES, 0, row0, 0, col0 >> 10, (col0 & 1023) >> 5);
          tma_store_commit();
        }
        buf ^= 1;
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
  if (warp == 2) { tc_fence_after(); tmem_dealloc<2 * BN>(tmem_base); }
}

// explicit instantiations used by api.cu
template __global__ void gemm_i8_i32_kernel<256>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip, const int);
template __global__ void gemm_i8_i32_kernel<128>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip, const int);
template __global__ void gemm_i8_i32_kernel<64>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip, const int);

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
