// gemm_i8_epilogue.cuh — the epilogue of the int8 projection GEMMs (gemm_i8.cu, gemm_i8_pair.cu): one warp's share of a
// 128 x 256 int32 accumulator, TMEM -> registers -> swizzled shared-memory staging -> TMA store.
//
// A chunk (32 rows x 32 columns = 4 KB) is a serial chain of long-latency steps — tcgen05.ld, wait::ld, the staging
// buffer's previous TMA store must have been READ, st.shared, proxy fence, TMA store — measured at ~ 900 cycles per chunk
// for a warp working alone, and longer when its stores queue behind the operand loads of the main loop in the SM's TMA
// unit.  With four warps that chain, not a bandwidth, set the tile time (stores only 7.0 K cycles, main loop only 6.0 K,
// both 10.9 K per tile: profiles/r02/gemm_attrib.txt).  Hence GEMM_EPI_WARPS = 8: two warps per TMEM lane quarter, each
// taking every other chunk, two staging buffers per warp.
#pragma once
#include "kernels.cuh"

namespace rnnt {

// t_row: TMEM address of (lane quarter, accumulator stage); cbuf0: this warp's two staging buffers; chunks
// c = first, first + step, ... < nchunks of the tile; (row0, col_base): where the warp's 32 rows x 256 columns go in C.
__device__ __forceinline__ void gemm_i8_store_chunks(const uint32_t t_row, const uint32_t cbuf0, int& buf, const int first,
                                                     const int step, const int nchunks, const CUtensorMap* tmap_c, const int gate_blocked,
                                                     const int row0, const int col_base, const int lane, const bool no_store) {
  constexpr uint32_t CB = 32 * 128;       // bytes per staging buffer
#pragma unroll 1
  for (int c = first; c < nchunks; c += step) {
    uint32_t v[32];
    RNNT_TMEM_LD_X16(t_row + c * 32, v);
    RNNT_TMEM_LD_X16(t_row + c * 32 + 16, v + 16);
    tmem_ld_wait();
    if (no_store) continue;
    // staging buffer `buf` was last read by the TMA store issued two chunks ago
    if (lane == 0) tma_store_wait_read<1>();
    __syncwarp();
    if (!gate_blocked) {
      const uint32_t dst = cbuf0 + (uint32_t)buf * CB + (uint32_t)lane * 128u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t a = dst + (uint32_t)((j ^ (lane & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v[4 * j]), "r"(v[4 * j + 1]),
                     "r"(v[4 * j + 2]), "r"(v[4 * j + 3]) : "memory");
      }
    } else if (gate_blocked == 2) {
      // gate-blocked result, M % 32 == 0 (see api.cu): the 32 rows x 8 units of an octet are ONE contiguous KB of C, and
      // the tensor map says so (inner dimension 256 int32, no swizzle) — the TMA unit then writes full 128-byte lines
      // with one L2 request per line; as 32-byte rows it issued one request per SECTOR and the L2 tag pipeline, shared
      // with the operand loads, was the busiest unit of the kernel (ncu round 2: lts__t_tag_requests 68 %, 282 M write
      // requests for 282 M sectors).  Staging = [octet 4][row 32][8 units], unswizzled; lanes 4-7 of every 8-lane store
      // phase write the other 16-byte half so that the phase covers all 32 banks.
      const uint32_t dst = cbuf0 + (uint32_t)buf * CB + (uint32_t)lane * 32u;
      const bool hs = ((lane >> 2) & 1) != 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int o = j >> 1;
        const bool hi = ((j & 1) != 0) != hs;           // which half of the octet this lane stores in this instruction
        const uint32_t x0 = hi ? v[8 * o + 4] : v[8 * o], x1 = hi ? v[8 * o + 5] : v[8 * o + 1];
        const uint32_t x2 = hi ? v[8 * o + 6] : v[8 * o + 2], x3 = hi ? v[8 * o + 7] : v[8 * o + 3];
        const uint32_t a = dst + (uint32_t)(o * 1024) + (hi ? 16u : 0u);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x0), "r"(x1), "r"(x2), "r"(x3) : "memory");
      }
    } else {
      // gate-blocked result, any M: rows of 32 bytes; staging = [unit octet 4][row 32][8 units] with the 32-byte TMA
      // swizzle (16-byte half ^= bit 2 of the row), which spreads the eight lanes of a store phase over all 32 banks
      const uint32_t dst = cbuf0 + (uint32_t)buf * CB + (uint32_t)lane * 32u;
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
      const int col0 = col_base + c * 32;
      if (!gate_blocked) tma_store_2d(tmap_c, cbuf0 + (uint32_t)buf * CB, col0, row0);
      else tma_store_5d(tmap_c, cbuf0 + (uint32_t)buf * CB, 0, gate_blocked == 2 ? row0 >> 5 : row0, 0, col0 >> 10, (col0 & 1023) >> 5);
      tma_store_commit();
    }
    buf ^= 1;
  }
}

}  // namespace rnnt
