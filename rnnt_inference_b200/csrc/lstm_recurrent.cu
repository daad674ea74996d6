// lstm_recurrent.cu — persistent recurrent kernel for one int8 LSTM layer (H = 1024) on sm_100a.
//
// Replaces phase B of lstm_layer_amx_int8 (mlperf_plugins/csrc/lstm_amx.cpp:122-152): per timestep the
// reference runs an AMX GEMM h.W_hh^T with the i8o32b32_accum epilogue (tile_product_tpp.hpp:308-345) and then
// lstm_postop_tpp::ref per row (tpps/lstm_postop_tpp.cpp:12-98) — two OpenMP regions and a round trip of the
// fp32 gates through memory per step.  Here all T steps run inside ONE cooperative launch:
//
//   grid = 32 hidden slices x 4 batch-tile groups (optionally clusters of CS slice-CTAs); CTA (s, y) owns hidden
//   units [32 s, 32 s + 32) of all four gates and walks the batch tiles y, y + 4, ... of 128 rows each.
//   * its 128 rows of W_hh (128 KB int8) are loaded ONCE by TMA and stay in shared memory for the whole sequence.
//     (A tcgen05.mma M128 x N x K32 costs max(64, N/2) cycles — the A-operand fetch — so N = 128 does twice the work of
//     N = 64 per h byte that passes through shared memory, the measured bound of the 16-unit version; profiles/r01.)
//   * per work item (t, batch tile):
//       - warp 0 (producer): TMA-loads the 128x1024 int8 h_t tile as eight 16 KB K-chunks through a 3-stage ring
//         (h lives in a K-chunk-major ping-pong buffer, so a chunk is one contiguous 16 KB read);
//       - warp 1: one thread issues tcgen05.mma.kind::i8 (M=128, N=128, K=32) x 32 into a double-buffered TMEM
//         accumulator;
//       - 16 epilogue warps: thread (row, 8 units) fetches its part of the input projection G[t] straight from global
//         memory (one 256-bit load = one full sector per gate) before waiting for the accumulator, then tcgen05.ld,
//         add G (int32, exact), dequantise, evaluate the fp16 gates, update c (kept between steps in a CTA-private
//         blocked scratch, so its traffic is fully coalesced), stage h_{t+1} (int8) and y[t] in shared memory; 4 of
//         them then write whole 32/64-byte row segments out;
//       - warp 3 (publisher): when the item's stores are issued, one release increment of the tile's counter.
//   * the only cross-CTA dependency is "h_t of batch tile b is complete": that per-tile counter (acquire by the
//     producers).  No grid-wide barrier.  Several batch tiles per CTA are in flight, so the publish -> poll -> load
//     latency of one tile is covered by the work of the others.
//   * ragged batches: tile b runs tile_steps[b] timesteps only (the SUT sorts utterances by length, so a tile's
//     rows have similar lengths); dead (t, tile) items are skipped by every role.
// See DESIGN.md for the byte/instruction budget and the measured timeline.
#include "kernels.cuh"
#include "gate_math.cuh"
#include "../../include/rnnt_rcp_ph_lut.h"

namespace rnnt {

__constant__ uint16_t c_rcp_lut[1024] = {RNNT_RCP_PH_LUT_VALUES};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}

// Work items of a CTA in issue order: for t, for its tiles, skipping tiles that have finished.  Start with (t=0, i=-1).
struct ItemCursor {
  int t, i;
  __device__ __forceinline__ bool next(int T, int my_tiles, const int* tiles, const int* steps) {
    if (my_tiles <= 0) return false;          // a tile group the host left empty
    for (;;) {
      if (++i >= my_tiles) { i = 0; if (++t >= T) return false; }
      if (t < steps[tiles[i]]) return true;
    }
  }
};

__device__ __forceinline__ void ldg_units_stream(const int32_t* ptr, int (&v)[8]) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.s32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(ptr));
}
__device__ __forceinline__ void ldg_units_stream(const int32_t* ptr, int (&v)[4]) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "l"(ptr));
}

template <int CS, int HS>
__global__ void __launch_bounds__(REC_THREADS, 1)
lstm_recurrent_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_h0,
                      const __grid_constant__ CUtensorMap tmap_h, const RecParams p) {
  using S = RecSmemT<HS>;
  constexpr int NS = S::NS, CPS = S::CPS;
  constexpr int UPT = S::UPT;              // hidden units per epilogue thread (8 | 4), processed as pairs
  constexpr int SPC = 128 / HS;            // slices per 128-byte K chunk of the h tile (4 | 8)
  constexpr uint16_t MC_MASK = (uint16_t)((1u << CS) - 1u);
  // 1024-byte alignment is declared (the dynamic window starts on such a boundary) rather than fixed up at run time, so
  // every shared-memory address below is base-constant + offset and folds into the instructions' immediate fields.
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  uint8_t* smem_gen = smem_raw;
  if (smem_base & 1023u) {          // never observed; uniform over the grid, so every CTA leaves before any barrier
    if (threadIdx.x == 0) atomicCAS(p.status, DEV_OK, DEV_BAD_SMEM);
    return;
  }
  const uint32_t smem_w = smem_base;
  const uint32_t smem_h = smem_base + S::W_BYTES;
  const uint32_t smem_out = smem_base + S::OUT_OFFSET;
  uint16_t* lut = reinterpret_cast<uint16_t*>(smem_gen + S::LUT_OFFSET);
  const uint32_t tab = smem_base + S::TAB_OFFSET;
  const uint32_t bar_base = smem_base + S::BAR_OFFSET;
  const uint32_t full_bar = bar_base;                  // [NS] h chunk landed (TMA complete_tx)
  const uint32_t empty_bar = bar_base + 64;            // [NS] h chunk consumed by the MMAs of all CS CTAs
  const uint32_t outfree_bar = bar_base + 128;         // staging buffer read by the storing warps (one item deep)
  const uint32_t tfull_bar = bar_base + 144;           // [2] accumulator complete
  const uint32_t done_bar = bar_base + 160;            // [2] item finished by all epilogue warps (frees the accumulator)
  const uint32_t staged_bar = bar_base + 176;          // [2] h / y rows of the item staged in shared memory
  const uint32_t wfull_bar = bar_base + 192;
  const uint32_t tmem_slot = bar_base + 200;
  int* s_steps = reinterpret_cast<int*>(smem_gen + S::BAR_OFFSET + 208);   // [REC_MAX_TILES]
  int* s_tile = s_steps + REC_MAX_TILES;                                   // [REC_MAX_TILES] tile ids of this group
  float* s_bias = reinterpret_cast<float*>(smem_gen + S::BIAS_OFFSET);     // [gate][unit of the slice]
  static_assert(NS * 8 <= 64 && 208 + 2 * REC_MAX_TILES * 4 <= 512, "barrier layout");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slice = blockIdx.x;
  const int num_slices = gridDim.x;
  // batch tiles of this CTA's tile group, dealt by the host (api.cu: the group that holds the longest tile gets the
  // shortest ones as company, so that no group is left alone with a long single-tile tail)
  const int by = blockIdx.y;
  const int my_tiles = p.group_count[by];
  const uint32_t cta_rank = (CS > 1) ? cluster_ctarank() : 0u;
  long long* const trace = RNNT_TRACE_PTR(p);
  const bool tracer = trace && blockIdx.x == 0 && blockIdx.y == 0;

  for (int i = threadIdx.x; i < 1024; i += REC_THREADS) lut[i] = c_rcp_lut[i];
  if (threadIdx.x < 4 * HS) s_bias[threadIdx.x] = p.bias[(threadIdx.x / HS) * REC_H + blockIdx.x * HS + (threadIdx.x % HS)];
  if (threadIdx.x < REC_MAX_TILES) {
    s_steps[threadIdx.x] = (threadIdx.x < p.num_btiles) ? p.tile_steps[threadIdx.x] : 0;
    s_tile[threadIdx.x] = p.group_tile[blockIdx.y][threadIdx.x];
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_w); tma_prefetch_desc(&tmap_h0); tma_prefetch_desc(&tmap_h);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NS; ++i) { mbar_init(full_bar + 8 * i, 1); mbar_init(empty_bar + 8 * i, CS); }
    mbar_init(outfree_bar, 2);
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull_bar + 8 * i, 1);
      mbar_init(done_bar + 8 * i, REC_EPI_WARPS);
      mbar_init(staged_bar + 8 * i, REC_EPI_WARPS);
    }
    mbar_init(wfull_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<2 * S::UMMA_N>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  {   // activation table: every reachable |x| evaluated once with the reference's instruction sequence
    const HalfConsts kc = make_half_consts();
    uint16_t* tabp = reinterpret_cast<uint16_t*>(smem_gen + S::TAB_OFFSET);
    for (int i = threadIdx.x * 2; i < REC_TAB_N; i += REC_THREADS * 2) {
      const __half2 r = pade_clamped(bits_h2((uint32_t)i | ((uint32_t)(i + 1) << 16)), kc, lut);
      *reinterpret_cast<uint32_t*>(tabp + i) = h2_bits(r);
    }
  }
  __syncthreads();
  if (CS > 1) cluster_sync_all();     // peers' barriers are initialised before any multicast / remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *(volatile uint32_t*)(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      mbar_expect_tx(wfull_bar, S::W_BYTES);
      for (int kc = 0; kc < REC_KC; ++kc)
        tma_load_3d(smem_w + kc * S::W_CHUNK, &tmap_w, wfull_bar, kc * 128, slice * HS, 0);
      bool ok = true;
      uint32_t chunk = 0;
      ItemCursor cur{0, -1};
      bool have = cur.next(p.T, my_tiles, s_tile, s_steps);
      while (have && ok) {
        const int t = cur.t, bt = s_tile[cur.i];
        if (t > 0 && !RNNT_DBG(p, 256)) {
          // wait until every slice CTA has published its part of h_t for this batch tile
          const int* flag = p.flags + bt * REC_FLAG_STRIDE;
          const int target = t * num_slices * 2;      // two storing warps per slice CTA
          long long t0 = clock64(); uint32_t spins = 0;
          for (;;) {
            int v;
            // relaxed polls (an acquire load costs an L1 invalidation per poll); one acquire fence after the loop
            asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
            if (v >= target) break;
            if ((++spins & 63u) == 0) {
              if (*(volatile int*)p.status != DEV_OK) { ok = false; break; }
              if (clock64() - t0 > kWatchdogCycles) { watchdog_fired(p.status, DEV_TIMEOUT_FLAG, 3, t); ok = false; break; }
            }
          }
          if (!ok) break;
          asm volatile("fence.acquire.gpu;" ::: "memory");
          if (!RNNT_DBG(p, 32)) fence_proxy_async_all();   // generic-proxy writes of h_t (other CTAs) -> async-proxy (TMA) reads
        }
        if (tracer && cur.i == 0) trace[t * 16 + 0] = clock64();
        const int hrow0 = (((t & 1) * p.num_btiles + bt) * REC_KC) * REC_BM;   // row of chunk 0 in the chunk-major h buffer
        for (int kd = 0; kd < REC_KC / CPS; ++kd, ++chunk) {          // `chunk` counts ring slots (CPS K-chunks each)
          const uint32_t st = chunk % NS, ph = (chunk / NS) & 1u;
          // the slot is free in EVERY CTA of the cluster once this barrier flips (multicast commits)
          if (!mbar_wait(empty_bar + 8 * st, ph ^ 1u, p.status)) { ok = false; break; }
          if (RNNT_DBG(p, 4)) { mbar_arrive(full_bar + 8 * st); continue; }
          mbar_expect_tx(full_bar + 8 * st, S::SLOT);
          const uint32_t dst = smem_h + st * S::SLOT, bar = full_bar + 8 * st;
          if (t > 0 && CS == 1) {
            // the ping-pong buffer holds every 16 KB chunk in the 128-byte-swizzled order the UMMA descriptor expects (the
            // storing warps write it that way) and the chunks of a tile follow each other, so a slot is ONE contiguous bulk
            // copy; clusters multicast chunk by chunk through a tensor map without swizzle
            bulk_load_1d(dst, p.h_buf + (size_t)(hrow0 + kd * CPS * REC_BM) * 128, S::SLOT, bar);
            continue;
          }
#pragma unroll
          for (int c = 0; c < CPS; ++c) {
            const int kc = kd * CPS + c;
            if (CS > 1 && (uint32_t)(kc % CS) != cta_rank) continue;
            if (t == 0) {            // h_0 straight from the caller's [N][1024] array (rows >= N read as zero)
              if (CS == 1) tma_load_2d(dst + c * S::H_CHUNK, &tmap_h0, bar, kc * 128, bt * REC_BM);
              else tma_load_2d_mc(dst + c * S::H_CHUNK, &tmap_h0, bar, kc * 128, bt * REC_BM, MC_MASK);
            } else {
              tma_load_2d_mc(dst + c * S::H_CHUNK, &tmap_h, bar, 0, hrow0 + kc * REC_BM, MC_MASK);
            }
          }
        }
        have = cur.next(p.T, my_tiles, s_tile, s_steps);
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_i8(REC_BM, S::UMMA_N);
      bool ok = mbar_wait(wfull_bar, 0, p.status);
      uint32_t chunk = 0; int acc = 0; uint32_t acc_phase = 0;
      ItemCursor cur{0, -1};
      while (ok && cur.next(p.T, my_tiles, s_tile, s_steps)) {
        const bool mtr = tracer && cur.i == 0;
        long long c0 = mtr ? clock64() : 0, w_full = 0;
        if (!mbar_wait(done_bar + 8 * acc, acc_phase ^ 1u, p.status)) break;
        if (mtr) trace[cur.t * 16 + 8] = clock64() - c0;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * S::UMMA_N;
        for (int kd = 0; kd < REC_KC / CPS; ++kd, ++chunk) {
          const uint32_t st = chunk % NS, ph = (chunk / NS) & 1u;
          if (mtr) c0 = clock64();
          if (!mbar_wait(full_bar + 8 * st, ph, p.status)) { ok = false; break; }
          if (mtr) w_full += clock64() - c0;
          if (kd == 0 && tracer && cur.i == 0) trace[cur.t * 16 + 1] = clock64();
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < CPS; ++c) {
            const int kc = kd * CPS + c;
            const uint64_t adesc = umma_desc_sw128(smem_h + st * S::SLOT + c * S::H_CHUNK);
            const uint64_t bdesc = umma_desc_sw128(smem_w + kc * S::W_CHUNK);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (!RNNT_DBG(p, 1)) umma_i8(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kc | k) != 0);
          }
          if (CS == 1) umma_commit(empty_bar + 8 * st);
          else umma_commit_mc(empty_bar + 8 * st, MC_MASK);
        }
        if (!ok) break;
        umma_commit(tfull_bar + 8 * acc);
        if (tracer && cur.i == 0) { trace[cur.t * 16 + 2] = clock64(); trace[cur.t * 16 + 9] = w_full; }
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (warp == 2 || warp == 3) {
    // ---------------------------------------------------------------- store + publish: each of the two warps writes 64
    // rows of the staged item (whole 32 / 64-byte row segments) and then publishes them with ONE release increment of
    // the tile's counter.  Keeping the global stores and their gpu-scope fence off the epilogue warps takes ~1.4k
    // cycles (the drain of the stores) out of every item's critical path.
    const int sw = warp - 2;
    int acc = 0; uint32_t acc_phase = 0;
    bool ok = true;
    ItemCursor cur{0, -1};
    while (ok && cur.next(p.T, my_tiles, s_tile, s_steps)) {
      const int t = cur.t, bt = s_tile[cur.i];
      const bool last = t + 1 == s_steps[bt];
      const bool tr = tracer && cur.i == 0 && sw == 0 && lane == 0;
      if (!mbar_wait(staged_bar + 8 * acc, acc_phase, p.status)) break;
      const uint32_t out_h = smem_out, out_y = out_h + S::OUT_H;
      constexpr int QUADS = HS / 4;          // unit quads of the slice = 32-bit words of int8 h per row (8 | 4)
      uint32_t hw[2][QUADS], yw[2][2 * QUADS];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int row = sw * 64 + r * 32 + lane;
#pragma unroll
        for (int i = 0; i < QUADS; ++i) hw[r][i] = lds32(out_h + (uint32_t)(i * REC_BM + row) * 4u);
        if (p.skip_quant_y) {
#pragma unroll
          for (int i = 0; i < QUADS; ++i) { const uint2 v = lds64(out_y + (uint32_t)(i * REC_BM + row) * 8u); yw[r][2 * i] = v.x; yw[r][2 * i + 1] = v.y; }
        } else {
#pragma unroll
          for (int i = 0; i < QUADS; ++i) yw[r][i] = lds32(out_y + (uint32_t)(i * REC_BM + row) * 4u);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(outfree_bar);          // staging buffer may be overwritten by the next item
      // h first: it is all the next step waits for, so only its stores are in flight when the release fence drains them
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int row = sw * 64 + r * 32 + lane;
        const int n = bt * REC_BM + row;
        // A tile that runs a single step keeps its final h in the ping-pong buffer too: other slice CTAs may still be
        // reading h_0 from hx (nothing orders t = 0 across CTAs); rec_copy_h1_kernel moves it to hx after the launch.
        if (!last || t == 0) {
          // 16-byte unit u of a row goes to position u ^ (row & 7): the shared-memory image of SWIZZLE_128B
          uint4* hn = reinterpret_cast<uint4*>(p.h_buf + ((((size_t)((t + 1) & 1) * p.num_btiles + bt) * REC_KC + slice / SPC) * REC_BM + row) * 128);
          const int u0 = (slice % SPC) * (HS / 16);
#pragma unroll
          for (int i = 0; i < QUADS / 4; ++i) hn[(u0 + i) ^ (row & 7)] = make_uint4(hw[r][4 * i], hw[r][4 * i + 1], hw[r][4 * i + 2], hw[r][4 * i + 3]);
        } else if (n < p.N) {   // last step of this tile: hand the state back to the caller's hx
          uint4* hp = reinterpret_cast<uint4*>(p.hx + (size_t)n * REC_H + slice * HS);
#pragma unroll
          for (int i = 0; i < QUADS / 4; ++i) hp[i] = make_uint4(hw[r][4 * i], hw[r][4 * i + 1], hw[r][4 * i + 2], hw[r][4 * i + 3]);
        }
      }
      if (tr) trace[t * 16 + 5] = clock64();
      __syncwarp();
      // h_{t+1}[tile bt][this slice][these 64 rows] is written: one release increment of the tile counter (the fence is
      // cumulative over the other lanes' stores through the warp barrier above)
      if (lane == 0 && !last) {
        if (tr) trace[t * 16 + 6] = clock64();
        if (RNNT_DBG(p, 16)) asm volatile("red.relaxed.gpu.global.add.s32 [%0], 1;" ::"l"(p.flags + bt * REC_FLAG_STRIDE) : "memory");
        else asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(p.flags + bt * REC_FLAG_STRIDE) : "memory");
        if (tr) trace[t * 16 + 7] = clock64();
      }
      __syncwarp();
      // y[t] afterwards: nothing inside this kernel reads it
      if (!RNNT_DBG(p, 8)) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int n = bt * REC_BM + sw * 64 + r * 32 + lane;
          if (n >= p.N) continue;
          const size_t yoff = ((size_t)t * p.N + n) * REC_H + slice * HS;
          if (p.skip_quant_y) {
            uint4* yp = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.y) + yoff);
#pragma unroll
            for (int i = 0; i < QUADS / 2; ++i) yp[i] = make_uint4(yw[r][4 * i], yw[r][4 * i + 1], yw[r][4 * i + 2], yw[r][4 * i + 3]);
          } else {
            uint4* yp = reinterpret_cast<uint4*>(reinterpret_cast<int8_t*>(p.y) + yoff);
#pragma unroll
            for (int i = 0; i < QUADS / 4; ++i) yp[i] = make_uint4(yw[r][4 * i], yw[r][4 * i + 1], yw[r][4 * i + 2], yw[r][4 * i + 3]);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp >= REC_CTRL_WARPS) {
    // ---------------------------------------------------------------- epilogue: gates, cell update, outputs
    const int e = warp - REC_CTRL_WARPS;
    const int q = warp & 3;                 // TMEM lane quarter accessible to this warp (warp id % 4)
    const int ug = e >> 2;                  // unit group within the slice (UPT = 8 | 4 units each, processed as quads)
    const int row = q * 32 + lane;          // batch row within the tile
    const int j0 = slice * HS + ug * UPT;   // first hidden unit of this thread
    const HalfConsts k = make_half_consts();
    const __half2 in_s = __float2half2_rn(p.in_scale), out_s = __float2half2_rn(p.out_scale);
    const bool same_scale = h2_bits(in_s) == h2_bits(out_s);
    const uint32_t bias_s = smem_base + S::BIAS_OFFSET + (uint32_t)(ug * UPT) * 4u;

    // c of an item: t == 0 from the caller's cx, else from the blocked scratch this same thread wrote one step earlier.
    // It is fetched one item ahead (cn) so that its L2 latency hides behind the previous item's gate math; when the next
    // item is the same tile (a CTA with one live tile) the freshly computed value is carried in registers instead.
    struct CVec { uint32_t w[UPT / 2]; };      // UPT binary16 values: one 16- or 8-byte access
    auto c_scratch = [&](int bt_) {
      return p.c_ws + ((((size_t)bt_ * (REC_H / HS) + slice) * 4 + ug) * REC_BM + row) * UPT;
    };
    auto ld_c = [&](const uint16_t* src, CVec& c) {
      if constexpr (UPT == 8) { const uint4 v = *reinterpret_cast<const uint4*>(src); c.w[0] = v.x; c.w[1] = v.y; c.w[2] = v.z; c.w[3] = v.w; }
      else { const uint2 v = *reinterpret_cast<const uint2*>(src); c.w[0] = v.x; c.w[1] = v.y; }
    };
    auto st_c = [&](uint16_t* dst, const CVec& c) {
      if constexpr (UPT == 8) *reinterpret_cast<uint4*>(dst) = make_uint4(c.w[0], c.w[1], c.w[2], c.w[3]);
      else *reinterpret_cast<uint2*>(dst) = make_uint2(c.w[0], c.w[1]);
    };
    auto load_c = [&](int t_, int bt_, CVec& c) {
      if (t_ == 0) {
        const int n_ = bt_ * REC_BM + row;
        if (n_ < p.N) ld_c(p.cx + (size_t)n_ * REC_H + j0, c);
        else {
#pragma unroll
          for (int i = 0; i < UPT / 2; ++i) c.w[i] = 0u;
        }
      } else {
        ld_c(c_scratch(bt_), c);
      }
    };
    // G of an item (this row's input projection, one full sector per gate) does not depend on the recurrent product: it
    // is requested before the wait for the accumulator.  (Fetching it a whole item ahead was measured slower.)
    int gi[4][UPT];
    auto load_g = [&](int t_, int bt_) {
      const int n_ = bt_ * REC_BM + row;
      if (n_ < p.N && !RNNT_DBG(p, 64)) {
        if (p.g_blocked) {      // [32-unit slice][gate][unit octet][row][8]: a warp's 32 rows are 1 KB contiguous per octet
          const size_t rows = (size_t)p.T * p.N;
          const int32_t* gp = p.g + (((size_t)(j0 >> 5) * 16 + ((j0 & 31) >> 3)) * rows + (size_t)t_ * p.N + n_) * 8 + (j0 & 7);
#pragma unroll
          for (int g = 0; g < 4; ++g) ldg_units_stream(gp + (size_t)g * 4 * rows * 8, gi[g]);
        } else {
          const int32_t* gp = p.g + ((size_t)t_ * p.N + n_) * (4 * REC_H) + j0;
#pragma unroll
          for (int g = 0; g < 4; ++g) ldg_units_stream(gp + g * REC_H, gi[g]);
        }
      } else {
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
          for (int u = 0; u < UPT; ++u) gi[g][u] = 0;
      }
    };
    int acc = 0; uint32_t acc_phase = 0, item = 0;
    ItemCursor cur{0, -1}, nxt{0, -1};
    bool have = cur.next(p.T, my_tiles, s_tile, s_steps);
    bool have_next = nxt.next(p.T, my_tiles, s_tile, s_steps);     // nxt runs one item ahead of cur
    if (have_next) have_next = nxt.next(p.T, my_tiles, s_tile, s_steps);
    CVec cn;
#pragma unroll
    for (int i = 0; i < UPT / 2; ++i) cn.w[i] = 0u;
    if (have) load_c(cur.t, s_tile[cur.i], cn);
    for (; have; have = cur.next(p.T, my_tiles, s_tile, s_steps), ++item) {
      const int t = cur.t, bt = s_tile[cur.i];
      const int steps = s_steps[bt];
      const int n = bt * REC_BM + row;
      const bool valid = n < p.N;
      uint32_t cbits[UPT / 2];
#pragma unroll
      for (int i = 0; i < UPT / 2; ++i) cbits[i] = cn.w[i];
      const bool carry = have_next && (s_tile[nxt.i]) == bt;       // next item = next step of this very tile
      if (have_next && !carry) load_c(nxt.t, s_tile[nxt.i], cn);
      const bool tr = tracer && cur.i == 0 && e == 0 && lane == 0;
      load_g(t, bt);
      const long long f0 = tr ? clock64() : 0;
      if (!mbar_wait(tfull_bar + 8 * acc, acc_phase, p.status)) break;
      if (tr) { trace[t * 16 + 3] = clock64(); trace[t * 16 + 11] = clock64() - f0; }
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * S::UMMA_N + ug * UPT);
      uint32_t hq_pack[UPT / 4], yq_pack[UPT / 4];
#pragma unroll
      for (int i = 0; i < UPT / 4; ++i) { hq_pack[i] = 0u; yq_pack[i] = 0u; }
      uint32_t ybf[UPT / 2];
      // all units of the thread at once: independent (pair) chains keep the table-lookup latency covered
      __half2 gate[4][UPT / 2];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint32_t a[UPT];
        if constexpr (UPT == 8)
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                       : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7])
                       : "r"(t_row + g * HS) : "memory");
        else
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "r"(t_row + g * HS) : "memory");
        tmem_ld_wait();
#pragma unroll
        for (int pr = 0; pr < UPT / 2; ++pr) {
          const uint2 bb = lds64(bias_s + (uint32_t)(g * HS + 2 * pr) * 4u);
          gate[g][pr] = __floats2half2_rn(dequant_gate((int)a[2 * pr], gi[g][2 * pr], __uint_as_float(bb.x), p.rb_scale),
                                          dequant_gate((int)a[2 * pr + 1], gi[g][2 * pr + 1], __uint_as_float(bb.y), p.rb_scale));
        }
      }
      tc_fence_before();
      if (tr) trace[t * 16 + 4] = clock64();
#pragma unroll
      for (int pr = 0; pr < UPT / 2; ++pr) {
        const int half = pr >> 1, u = (pr & 1) * 2;
        const __half2 it_ = sigmoid_tab2(gate[0][pr], k, tab);
        const __half2 ft_ = sigmoid_tab2(gate[1][pr], k, tab);
        const __half2 gt_ = tanh_tab2(gate[2][pr], tab);
        const __half2 ot_ = sigmoid_tab2(gate[3][pr], k, tab);
        const __half2 c_old = bits_h2(cbits[pr]);
        const __half2 c_new = __hfma2(it_, gt_, __hmul2_rn(ft_, c_old));     // c = fma(i, g, f*c)
        cbits[pr] = h2_bits(c_new);
        const __half2 ht = __hmul2_rn(ot_, tanh_tab2(c_new, tab));
        const uint32_t hq = quant_ph2(ht, in_s);
        hq_pack[half] |= hq << (8 * u);
        if (p.skip_quant_y) {
          const float2 hf = __half22float2(ht);
          ybf[pr] = float_to_bf16_bits(hf.x) | (float_to_bf16_bits(hf.y) << 16);
        } else {
          yq_pack[half] |= (same_scale ? hq : quant_ph2(ht, out_s)) << (8 * u);
        }
      }
      // ---- c: final value to the caller's array, otherwise to the CTA-private blocked scratch (coalesced 512 B per warp)
      const bool last = t + 1 == steps;
      CVec cv;
#pragma unroll
      for (int i = 0; i < UPT / 2; ++i) cv.w[i] = cbits[i];
      if (last) {
        if (valid) st_c(p.cx + (size_t)n * REC_H + j0, cv);
      } else if (carry) {
        cn = cv;
      } else {
        st_c(c_scratch(bt), cv);
      }
      // ---- stage h_{t+1} and y[t] as [unit quad][row] so that four warps can write whole 32 / 64-byte row segments.
      // The staging buffer is one item deep: wait until the storing warps have read the previous item out of it.
      if (!mbar_wait(outfree_bar, (item & 1u) ^ 1u, p.status)) break;
      const uint32_t out_h = smem_out, out_y = out_h + S::OUT_H;
#pragma unroll
      for (int half = 0; half < UPT / 4; ++half) {
        const uint32_t slot = (uint32_t)((ug * (UPT / 4) + half) * REC_BM + row);
        sts32(out_h + slot * 4u, hq_pack[half]);
        if (p.skip_quant_y) sts64(out_y + slot * 8u, ybf[half * 2], ybf[half * 2 + 1]);
        else sts32(out_y + slot * 4u, yq_pack[half]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(staged_bar + 8 * acc);
      // item finished by this warp.  A warp can be at most one item ahead of the slowest one (the next-but-one MMA waits
      // on done_bar), so barriers / counters indexed by item parity suffice.
      __syncwarp();
      if (lane == 0) mbar_arrive(done_bar + 8 * acc);
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      if (have_next) have_next = nxt.next(p.T, my_tiles, s_tile, s_steps);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CS > 1) cluster_sync_all();     // no CTA leaves while peers may still multicast into it / arrive on its barriers
  if (warp == 2) { tc_fence_after(); tmem_dealloc<2 * S::UMMA_N>(tmem_base); }
}

#define RNNT_REC_INST(CS, HS)                                                                                            \
  template __global__ void lstm_recurrent_kernel<CS, HS>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, \
                                                         const __grid_constant__ CUtensorMap, const RecParams);
RNNT_REC_INST(1, 32) RNNT_REC_INST(2, 32) RNNT_REC_INST(4, 32) RNNT_REC_INST(8, 32)
RNNT_REC_INST(1, 16) RNNT_REC_INST(2, 16) RNNT_REC_INST(4, 16) RNNT_REC_INST(8, 16)

}  // namespace rnnt
