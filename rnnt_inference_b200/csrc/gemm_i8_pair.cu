// gemm_i8_pair.cu — the int8 projection GEMM on CTA PAIRS (tcgen05 cta_group::2), for sm_100a.
//
// Same contract as gemm_i8.cu (C[M][4H] int32 = X[M][K] . W_ih[4H][K]^T, raw int32, row-major or gate-blocked, ragged
// tile skipping).  gemm_i8.cu is bound by the local shared-memory port (profiles/r02/gemm_attrib.txt): per 128 x 256
// tile every CTA stages 128 + 256 rows of operands and the UMMAs read them back.  Here two CTAs of a cluster (one TPC)
// compute ONE 256 x 256 tile with M256 x N256 x K32 MMAs issued by the leader CTA: each CTA stages its own 128 rows of A
// and only HALF of the W tile (128 rows), the tensor cores of both SMs read both halves — 128 + 128 rows per CTA and
// 128 x 256 outputs, i.e. a third less shared-memory traffic per output, and room for a 5-stage operand ring beside 64 KB of result staging.
//   warp 0 (both CTAs): TMA producer — its A box and its half of the W box into its own shared memory
//                       (cp.async.bulk.tensor .cta_group::2), completion bytes on the LEADER's full barrier
//   warp 1 (leader)   : one thread issues tcgen05.mma.cta_group::2.kind::i8; tcgen05.commit multicast frees the stage in
//                       both CTAs and publishes the accumulator to both epilogues
//   warp 2 (both)     : TMEM allocation (cta_group::2)
//   warps 4-11 (both) : epilogue of the CTA's own 128 rows (gemm_i8_epilogue.cuh), then an arrive
//                       on the leader's accumulator-free barrier (remote for the peer)
#include "gemm_i8_epilogue.cuh"

namespace rnnt {

__global__ void __launch_bounds__(GEMM_I8_THREADS, 1)
gemm_i8_i32_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                        const __grid_constant__ CUtensorMap tmap_c, int num_m_tiles, int num_n_tiles, int num_k_blocks,
                        int* __restrict__ status, const RowSkip skip, const int gate_blocked) {
  using S = GemmPairSmem;
  constexpr int BN = 256;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_base + S::STAGES * S::A_BYTES;
  const uint32_t smem_c = smem_base + S::STAGES * S::STAGE_BYTES;
  const uint32_t bar_base = smem_base + S::BAR_OFFSET;
  const uint32_t full_bar = bar_base;                         // [STAGES]  (the leader's are used)
  const uint32_t empty_bar = bar_base + 8 * S::STAGES;        // [STAGES]  one multicast commit per use, in both CTAs
  const uint32_t tfull_bar = bar_base + 16 * S::STAGES;       // [2]       accumulator complete, in both CTAs
  const uint32_t tempty_bar = tfull_bar + 16;                 // [2]       (the leader's: 4 + 4 epilogue warps)
  const uint32_t tmem_slot = tempty_bar + 16;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_m_super = (num_m_tiles + 1) >> 1;
  const int num_super = num_m_super * num_n_tiles;
  auto m_of = [&](int st, int r) { return (st / num_n_tiles) * 2 + r; };
  auto alive_at = [&](int st, int r) { const int m = m_of(st, r); return m < num_m_tiles && !skip.dead((long long)m * GEMM_BM); };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S::STAGES; ++i) { mbar_init(full_bar + 8 * i, 1); mbar_init(empty_bar + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull_bar + 8 * i, 1); mbar_init(tempty_bar + 8 * i, 2 * GEMM_I8_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_pair<2 * BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // both CTAs' barriers and TMEM allocations exist before anything crosses the pair
  tc_fence_after();
  const uint32_t tmem_base = *(volatile uint32_t*)(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; bool ok = true;
      for (int st = pair_id; st < num_super && ok; st += num_pairs) {
        if (!alive_at(st, 0) && !alive_at(st, 1)) continue;
        const int m_load = min(m_of(st, (int)rank), num_m_tiles - 1);      // a half past the matrix re-reads the last tile (result dropped)
        const int n_blk = st % num_n_tiles;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          if (!mbar_wait(empty_bar + 8 * stage, phase ^ 1, status)) { ok = false; break; }
          const uint32_t lead_full = map_to_cta(full_bar + 8 * stage, 0);
          if (leader) mbar_expect_tx(full_bar + 8 * stage, 2 * S::STAGE_BYTES);     // the bytes of both CTAs land on this barrier
          tma_load_2d_pair(smem_a + stage * S::A_BYTES, &tmap_a, lead_full, kb * GEMM_BK, m_load * GEMM_BM);
          tma_load_2d_pair(smem_b + stage * S::B_BYTES, &tmap_b, lead_full, kb * GEMM_BK, n_blk * BN + (int)rank * (BN / 2));
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (leader CTA only)
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = umma_idesc_i8(2 * GEMM_BM, BN);
      int stage = 0; uint32_t phase = 0; int acc = 0; uint32_t acc_phase = 0; bool ok = true;
      for (int st = pair_id; st < num_super && ok; st += num_pairs) {
        if (!alive_at(st, 0) && !alive_at(st, 1)) continue;
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
            umma_i8_pair(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          umma_commit_pair(empty_bar + 8 * stage, 0b11);     // frees the stage in both CTAs when these MMAs retire
          if (++stage == S::STAGES) { stage = 0; phase ^= 1; }
        }
        if (!ok) break;
        umma_commit_pair(tfull_bar + 8 * acc, 0b11);          // accumulator complete -> both epilogues
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue of this CTA's 128 rows
    const int q = warp & 3;
    const int half = (warp - 4) >> 2;             // the two warps of a TMEM lane quarter take alternate 32-column chunks
    const uint32_t cbuf0 = smem_c + (uint32_t)((warp - 4) * 2) * S::C_WARP_BYTES;
    const uint32_t lead_tempty = map_to_cta(tempty_bar, 0);
    int acc = 0; uint32_t acc_phase = 0; int buf = 0;
    for (int st = pair_id; st < num_super; st += num_pairs) {
      if (!alive_at(st, 0) && !alive_at(st, 1)) continue;
      const bool alive = alive_at(st, (int)rank);
      const int m_blk = m_of(st, (int)rank), n_blk = st % num_n_tiles;
      if (!mbar_wait(tfull_bar + 8 * acc, acc_phase, status)) break;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if (alive)
        gemm_i8_store_chunks(t_row, cbuf0, buf, half, GEMM_I8_EPI_WARPS / 4, BN / 32, &tmap_c, gate_blocked,
                             m_blk * GEMM_BM + q * 32, n_blk * BN, lane, false);
      // this warp is done with the accumulator (nothing to read for a dead half): tell the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(tempty_bar + 8 * acc);
        else mbar_arrive_cluster(lead_tempty + 8 * acc);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                 // neither CTA leaves (or frees TMEM) while the pair's MMAs / arrives may still touch it
  if (warp == 2) { tc_fence_after(); tmem_dealloc_pair<2 * BN>(tmem_base); }
}

}  // namespace rnnt
