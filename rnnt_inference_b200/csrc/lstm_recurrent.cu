// lstm_recurrent.cu — persistent recurrent kernel for one int8 LSTM layer (H = 1024) on sm_100a.
//
// Replaces phase B of lstm_layer_amx_int8 (mlperf_plugins/csrc/lstm_amx.cpp:122-152): per timestep the
// reference runs an AMX GEMM h.W_hh^T with the i8o32b32_accum epilogue (tile_product_tpp.hpp:308-345) and then
// lstm_postop_tpp::ref per row (tpps/lstm_postop_tpp.cpp:12-98) — two OpenMP regions and a round trip of the
// fp32 gates through memory per step.  Here all T steps run inside ONE cooperative launch:
//
//   grid = (1024/HS hidden slices) x (batch-tile groups); CTA (s, y) owns hidden units [s*HS, s*HS+HS) of all
//   four gates and walks the batch tiles y, y+gridDim.y, ... of 128 rows each.
//   * W_hh rows of the slice ([4*HS] x 1024 int8, 32/64 KB) are loaded ONCE by TMA and stay in shared memory.
//   * per (t, batch tile): TMA streams the 128x1024 int8 h tile through an 8-stage ring (one 128-byte
//     K-chunk per stage), one elected thread issues tcgen05.mma.kind::i8 (M=128, N=4*HS, K=32) into a
//     double-buffered TMEM accumulator, 16 epilogue warps read it back (tcgen05.ld), add the precomputed
//     input projection G[t] (int32, exact), dequantise, evaluate the fp16 gates, update c, and write
//     h_{t+1} (int8) to the ping-pong h buffer in L2 plus y[t].
//   * the only cross-CTA dependency is "h_t of batch tile b is complete": a per-tile arrival counter in
//     global memory (release by every slice CTA, acquire by the TMA producer).  No grid-wide barrier.
// Bounded by: SMEM reads of the two MMA operands (192 KB per tile-step at 128 B/clk), the fp16 gate math
// (~100 instructions per cell), and the L2 hand-off latency; see DESIGN.md.
#include "kernels.cuh"
#include "gate_math.cuh"
#include "../../include/rnnt_rcp_ph_lut.h"

namespace rnnt {


__constant__ uint16_t c_rcp_lut[1024] = {RNNT_RCP_PH_LUT_VALUES};


template <int HS>
__global__ void __launch_bounds__(REC_THREADS, 1)
lstm_recurrent_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_h,
                      const RecParams p) {
  using S = RecSmem<HS>;
  constexpr int UPT = HS / (REC_EPI_WARPS / 4);   // hidden units per epilogue thread (4 or 2)
  static_assert(UPT == 2 || UPT == 4, "HS must be 8 or 16");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t smem_w = smem_base;
  const uint32_t smem_h = smem_base + S::W_BYTES;
  uint16_t* lut = reinterpret_cast<uint16_t*>(smem_gen + S::LUT_OFFSET);
  const uint32_t bar_base = smem_base + S::BAR_OFFSET;
  const uint32_t full_bar = bar_base;                  // [8]
  const uint32_t empty_bar = bar_base + 64;            // [8]
  const uint32_t tfull_bar = bar_base + 128;           // [2]
  const uint32_t tempty_bar = bar_base + 144;          // [2]
  const uint32_t wfull_bar = bar_base + 160;
  const uint32_t epi_bar = bar_base + 168;             // epilogue warps -> publisher
  const uint32_t tmem_slot = bar_base + 176;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slice = blockIdx.x;
  const int num_slices = gridDim.x;
  const int my_tiles = (p.num_btiles - (int)blockIdx.y + (int)gridDim.y - 1) / (int)gridDim.y;
  const int num_items = p.T * my_tiles;

  for (int i = threadIdx.x; i < 1024; i += REC_THREADS) lut[i] = c_rcp_lut[i];
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmap_w); tma_prefetch_desc(&tmap_h); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < REC_KC; ++i) { mbar_init(full_bar + 8 * i, 1); mbar_init(empty_bar + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull_bar + 8 * i, 1); mbar_init(tempty_bar + 8 * i, REC_EPI_WARPS); }
    mbar_init(wfull_bar, 1);
    mbar_init(epi_bar, REC_EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<(2 * S::UMMA_N < 32 ? 32 : 2 * S::UMMA_N)>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *(volatile uint32_t*)(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      // resident weight slice: 8 chunks of [4 gates][HS units][128 B]
      mbar_expect_tx(wfull_bar, S::W_BYTES);
      for (int kc = 0; kc < REC_KC; ++kc)
        tma_load_3d(smem_w + kc * S::W_CHUNK, &tmap_w, wfull_bar, kc * 128, slice * HS, 0);
      uint32_t phase = 0; bool ok = true;
      for (int it = 0; it < num_items && ok; ++it) {
        const int t = it / my_tiles;
        const int bt = (int)blockIdx.y + (it % my_tiles) * (int)gridDim.y;
        if (t > 0) {
          // wait until every slice CTA has published its part of h_t for this batch tile
          const int* flag = p.flags + bt * REC_FLAG_STRIDE;
          const int target = t * num_slices;
          long long t0 = clock64(); uint32_t spins = 0;
          for (;;) {
            int v;
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
            if (v >= target) break;
            if ((++spins & 63u) == 0) {
              if (*(volatile int*)p.status != DEV_OK) { ok = false; break; }
              if (clock64() - t0 > kWatchdogCycles) { atomicCAS(p.status, DEV_OK, DEV_TIMEOUT_FLAG); ok = false; break; }
            }
          }
          if (!ok) break;
          fence_proxy_async_all();   // generic-proxy writes of h_t (other CTAs) -> async-proxy (TMA) reads
        }
        if (p.trace && blockIdx.x == 0 && blockIdx.y == 0) p.trace[t * 8 + 0] = clock64();
        for (int kc = 0; kc < REC_KC; ++kc) {
          if (!mbar_wait(empty_bar + 8 * kc, phase ^ 1, p.status)) { ok = false; break; }
          mbar_expect_tx(full_bar + 8 * kc, S::H_CHUNK);
          tma_load_3d(smem_h + kc * S::H_CHUNK, &tmap_h, full_bar + 8 * kc, kc * 128, bt * REC_BM, t & 1);
        }
        phase ^= 1;
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_i8(REC_BM, S::UMMA_N);
      bool ok = mbar_wait(wfull_bar, 0, p.status);
      uint32_t phase = 0; int acc = 0; uint32_t acc_phase = 0;
      for (int it = 0; it < num_items && ok; ++it) {
        if (!mbar_wait(tempty_bar + 8 * acc, acc_phase ^ 1, p.status)) break;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * S::UMMA_N;
        for (int kc = 0; kc < REC_KC; ++kc) {
          if (!mbar_wait(full_bar + 8 * kc, phase, p.status)) { ok = false; break; }
          if (kc == 0 && p.trace && blockIdx.x == 0 && blockIdx.y == 0) p.trace[(it / my_tiles) * 8 + 1] = clock64();
          tc_fence_after();
          const uint64_t adesc = umma_desc_sw128(smem_h + kc * S::H_CHUNK);
          const uint64_t bdesc = umma_desc_sw128(smem_w + kc * S::W_CHUNK);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_i8(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kc | k) != 0);
          umma_commit(empty_bar + 8 * kc);
        }
        if (!ok) break;
        umma_commit(tfull_bar + 8 * acc);
        if (p.trace && blockIdx.x == 0 && blockIdx.y == 0) p.trace[(it / my_tiles) * 8 + 2] = clock64();
        phase ^= 1;
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= REC_CTRL_WARPS) {
    // ---------------------------------------------------------------- epilogue: gates, cell update, outputs
    const int e = warp - REC_CTRL_WARPS;
    const int q = warp & 3;                 // TMEM lane quarter accessible to this warp (warp id % 4)
    const int ug = e >> 2;                  // unit group within the slice
    const int row = q * 32 + lane;          // batch row within the tile
    const int j0 = slice * HS + ug * UPT;   // first hidden unit of this thread
    const HalfConsts k = make_half_consts();
    const __half2 in_s = __float2half2_rn(p.in_scale), out_s = __float2half2_rn(p.out_scale);
    float bias[4][UPT];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int u = 0; u < UPT; ++u) bias[g][u] = p.bias[g * REC_H + j0 + u];

    int acc = 0; uint32_t acc_phase = 0, epi_phase = 0;
    for (int it = 0; it < num_items; ++it) {
      const int t = it / my_tiles;
      const int bt = (int)blockIdx.y + (it % my_tiles) * (int)gridDim.y;
      const int n = bt * REC_BM + row;
      const bool valid = n < p.N;
      // operands that do not depend on the recurrent product: fetch them before waiting on the MMA
      int gi[4][UPT];
      uint32_t cbits[UPT / 2];
      if (valid) {
        const int32_t* gp = p.G + ((size_t)t * p.N + n) * (4 * REC_H) + j0;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          if constexpr (UPT == 4) {
            const int4 v = __ldg(reinterpret_cast<const int4*>(gp + g * REC_H));
            gi[g][0] = v.x; gi[g][1] = v.y; gi[g][2] = v.z; gi[g][3] = v.w;
          } else {
            const int2 v = __ldg(reinterpret_cast<const int2*>(gp + g * REC_H));
            gi[g][0] = v.x; gi[g][1] = v.y;
          }
        }
        const uint32_t* cp = reinterpret_cast<const uint32_t*>(p.cx + (size_t)n * REC_H + j0);
#pragma unroll
        for (int u = 0; u < UPT / 2; ++u) cbits[u] = cp[u];
      } else {
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
          for (int u = 0; u < UPT; ++u) gi[g][u] = 0;
#pragma unroll
        for (int u = 0; u < UPT / 2; ++u) cbits[u] = 0;
      }

      if (!mbar_wait(tfull_bar + 8 * acc, acc_phase, p.status)) break;
      const bool tr = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && e == 0 && lane == 0;
      if (tr) p.trace[t * 8 + 3] = clock64();
      tc_fence_after();
      uint32_t a[4][UPT];
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * S::UMMA_N + ug * UPT);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        if constexpr (UPT == 4) {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(a[g][0]), "=r"(a[g][1]), "=r"(a[g][2]), "=r"(a[g][3])
                       : "r"(t_row + g * HS) : "memory");
        } else {
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];"
                       : "=r"(a[g][0]), "=r"(a[g][1]) : "r"(t_row + g * HS) : "memory");
        }
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);   // accumulator stage may be overwritten
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }

      if (tr) p.trace[t * 8 + 4] = clock64();
      uint32_t hq_pack = 0, yq_pack = 0;
      uint32_t ybf[UPT / 2];
#pragma unroll
      for (int u = 0; u < UPT; u += 2) {
        __half2 gate[4];
#pragma unroll
        for (int g = 0; g < 4; ++g)
          gate[g] = __floats2half2_rn(dequant_gate((int)a[g][u], gi[g][u], bias[g][u], p.rb_scale),
                                      dequant_gate((int)a[g][u + 1], gi[g][u + 1], bias[g][u + 1], p.rb_scale));
        const __half2 it_ = sigmoid_ph2(gate[0], k, lut);
        const __half2 ft_ = sigmoid_ph2(gate[1], k, lut);
        const __half2 gt_ = tanh_ph2(gate[2], k, lut);
        const __half2 ot_ = sigmoid_ph2(gate[3], k, lut);
        const __half2 c_old = bits_h2(cbits[u / 2]);
        const __half2 c_new = __hfma2(it_, gt_, __hmul2_rn(ft_, c_old));     // c = fma(i, g, f*c)
        cbits[u / 2] = h2_bits(c_new);
        const __half2 ht = __hmul2_rn(ot_, tanh_ph2(c_new, k, lut));
        hq_pack |= quant_ph2(ht, in_s) << (8 * u);
        if (p.skip_quant_y) {
          const float2 hf = __half22float2(ht);
          ybf[u / 2] = float_to_bf16_bits(hf.x) | (float_to_bf16_bits(hf.y) << 16);
        } else {
          yq_pack |= quant_ph2(ht, out_s) << (8 * u);
        }
      }
      if (valid) {
        int8_t* hn = p.h_buf + ((size_t)((t + 1) & 1) * p.N_pad + n) * REC_H + j0;
        if constexpr (UPT == 4) *reinterpret_cast<uint32_t*>(hn) = hq_pack;
        else *reinterpret_cast<uint16_t*>(hn) = (uint16_t)hq_pack;
        uint32_t* cp = reinterpret_cast<uint32_t*>(p.cx + (size_t)n * REC_H + j0);
#pragma unroll
        for (int u = 0; u < UPT / 2; ++u) cp[u] = cbits[u];
        const size_t yoff = ((size_t)t * p.N + n) * REC_H + j0;
        if (p.skip_quant_y) {
          uint32_t* yp = reinterpret_cast<uint32_t*>(reinterpret_cast<uint16_t*>(p.y) + yoff);
#pragma unroll
          for (int u = 0; u < UPT / 2; ++u) yp[u] = ybf[u];
        } else {
          int8_t* yp = reinterpret_cast<int8_t*>(p.y) + yoff;
          if constexpr (UPT == 4) *reinterpret_cast<uint32_t*>(yp) = yq_pack;
          else *reinterpret_cast<uint16_t*>(yp) = (uint16_t)yq_pack;
        }
      }
      // publish h_{t+1}[tile bt][this slice]: generic -> async proxy fence for the TMA readers, every
      // epilogue warp arrives on epi_bar (release.cta), the publisher thread waits for all 16 (bounded, so a
      // protocol error cannot hang a bar.sync), then does one release increment of the tile counter.
      if (tr) p.trace[t * 8 + 5] = clock64();
      fence_proxy_async_all();
      __syncwarp();
      if (tr) p.trace[t * 8 + 6] = clock64();
      if (lane == 0) mbar_arrive(epi_bar);
      if (e == 0 && lane == 0) {
        if (mbar_wait(epi_bar, epi_phase, p.status)) {
          __threadfence();
          asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(p.flags + bt * REC_FLAG_STRIDE) : "memory");
          if (tr) p.trace[t * 8 + 7] = clock64();
        }
      }
      epi_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) { tc_fence_after(); tmem_dealloc<(2 * S::UMMA_N < 32 ? 32 : 2 * S::UMMA_N)>(tmem_base); }
}

template __global__ void lstm_recurrent_kernel<16>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const RecParams);
template __global__ void lstm_recurrent_kernel<8>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, const RecParams);

}  // namespace rnnt
