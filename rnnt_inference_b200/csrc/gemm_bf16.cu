// gemm_bf16.cu — bf16 x bf16 -> fp32 (TMEM) -> bf16 GEMM for sm_100a: the hoisted transcription half of the joint fc1.
//
// The reference evaluates y1 = bf16(f . W_trans^T) inside every decode iteration with the AMX kernel i16o16b0
// (mlperf_plugins/csrc/amx_linear.cpp:137-150, tpps/tile_product_tpp.hpp:412-425) and ROUNDS IT TO bf16 IN MEMORY before
// the prediction half is accumulated on top — so y1 depends on the encoder frame only and can be computed once per
// frame for the whole batch: M = T2*N rows, K = 1024, N = 512 (2.9e11 FLOP per Offline query of 2513 utterances).
// A dense contraction of that size belongs on the tensor cores:
//   warp 0   : TMA producer — 128 x 64 bf16 (A) and BN x 64 bf16 (W) boxes, SWIZZLE_128B, 4-stage mbarrier ring
//   warp 1   : one thread issues tcgen05.mma.kind::f16 (BF16 x BF16 -> F32, M=128, N=BN, K=16) x4 per stage into a
//              double-buffered TMEM accumulator (2*BN <= 512 columns)
//   warp 2   : TMEM allocator
//   warps 4-7: epilogue — tcgen05.ld 32 lanes x 32 columns, fp32 -> bf16 (round to nearest even, like VCVTNE2PS2BF16),
//              64-byte rows staged with the 64-byte TMA swizzle (conflict-free stores), TMA store of a 32 x 32 box
// Accumulation order inside the tensor core differs from AMX's TDPBF16PS (and from the CPU oracle's); SURVEY 7.2 rules
// this tier "tolerance + top-2 margin": tests/test_amx_golden.py pins <= 1 bf16 ulp against real AMX output.
// Ragged batches: (timestep, 128-row tile) blocks past a tile's last step are skipped (RowSkip), like the int8 GEMM.
#include "kernels.cuh"

namespace rnnt {

// Instruction descriptor for kind::f16, BF16 x BF16 -> F32, both operands K-major
// (cute::UMMA::InstrDescriptor: c_format [4,6) = 1 (F32), a_format [7,10) = 1 (BF16), b_format [10,13) = 1, N>>3 [17,23), M>>4 [24,29)).
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));     // first source -> upper half
  return r;
}

template <int BN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles, int num_k_blocks,
                      int* __restrict__ status, const RowSkip skip) {
  using S = GemmBf16Smem<BN>;
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
          tma_load_2d(smem_a + stage * S::A_BYTES, &tmap_a, full_bar + 8 * stage, kb * S::BK_ELEMS, m_blk * GEMM_BM);
          tma_load_2d(smem_b + stage * S::B_BYTES, &tmap_b, full_bar + 8 * stage, kb * S::BK_ELEMS, n_blk * BN);
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
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
          for (int k = 0; k < S::BK_ELEMS / 16; ++k)        // K = 16 bf16 = 32 bytes per instruction: +2 in 16-byte units
            umma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          umma_commit(empty_bar + 8 * stage);   // frees the smem stage when these MMAs retire
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
        if (!ok) break;
        umma_commit(tfull_bar + 8 * acc);       // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue (TMEM -> bf16 -> smem -> TMA store)
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const uint32_t cbuf0 = smem_c + (uint32_t)(q * 2) * S::C_WARP_BYTES;
    const uint32_t sw = (uint32_t)((lane >> 1) & 3);          // 64-byte swizzle: 16-byte chunk ^= bits [1,3) of the row
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
        uint32_t o[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) o[j] = pack_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
        // staging buffer `buf` was last read by the TMA store issued two chunks ago
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        const uint32_t dst = cbuf0 + (uint32_t)buf * S::C_WARP_BYTES + (uint32_t)lane * 64u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t a = dst + ((((uint32_t)j) ^ sw) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(o[4 * j]), "r"(o[4 * j + 1]),
                       "r"(o[4 * j + 2]), "r"(o[4 * j + 3]) : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmap_c, cbuf0 + (uint32_t)buf * S::C_WARP_BYTES, n_blk * BN + c * 32, m_blk * GEMM_BM + q * 32);
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

template __global__ void gemm_bf16_bf16_kernel<256>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip);
template __global__ void gemm_bf16_bf16_kernel<128>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip);
template __global__ void gemm_bf16_bf16_kernel<64>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, int, int, int, int* __restrict__, const RowSkip);

}  // namespace rnnt
