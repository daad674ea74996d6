// greedy_decode.cu — the whole greedy decode loop as ONE device-resident cooperative kernel.
//
// Replaces GreedyDecoder.greedy_decode_quant's while-loop (models/decoder.py:188-211), in which every iteration
// dispatches Prediction (modeling_rnnt.py:183-205 -> lstm_amx_bf16, lstm_amx.cpp:160-266), Joint
// (:259-289 -> amx_linear_bf16_accum_relu + amx_linear_i16o32, amx_linear.cpp:113-170), torch.argmax over 32
// columns (decoder.py:193) and greedy_decode_update (greedy_decode_update.cpp:13-96) and returns a bool to Python.
//
// Arithmetic contract (matches oracle/rnnt_oracle.cpp): bf16 operands, fp32 FMA accumulation sequential in k
// from 0 (one thread owns one output chain — no cross-thread reduction), epilogues in the reference's order:
// G = (h.W_hh + x.W_ih) + bias; fp32 Pade sigmoid/tanh with RN reciprocal; y = bf16(relu((g.W_pred + y1) + b1));
// logits = y.W2 + b2; first-maximum argmax over 32 columns.  Hence token sequences are bit-exact vs the oracle.
//
// Organisation: the batch is split into NG lane groups; a group is served by S=20 CTAs that keep the
// prediction-LSTM weight columns of 16 hidden units (both layers, W_ih|W_hh, 162 KB) and their 26 rows of W_pred
// resident in shared memory.  Per iteration and group: (A) LSTM layer 0, (B) layer 1, (C) g.W_pred, (D) per lane:
// fc1 epilogue, fc2, argmax, state update — with a group-local barrier (arrival counter in global memory) between
// phases.  Groups never wait for each other inside the loop.  The loop is latency-bound (a handful of active lanes
// per group and iteration), so every phase issues all of its global loads up front (one L2 round trip), per-lane
// state travels as one 16-byte word, and inter-CTA data is read with ld.global.cg.  Ragged batch, per-utterance
// counters:
//   * the prediction network is recomputed only for lanes whose previous symbol was accepted: the reference
//     recomputes it every iteration, but with unchanged (pre_g, pre_hg, pre_cg) the result is bit-identical;
//   * a finished lane whose last step was a blank is at a fixed point of the reference loop and is skipped; a
//     lane that finished on the 30-symbol cap keeps emitting in the reference until the whole batch is done
//     (greedy_decode_update_tpp.cpp:22-29 masks only f_lens == 0), which is reproduced exactly: phase 1 runs
//     every group until its lanes finished, one grid-wide rendezvous yields the batch iteration count, phase 2
//     replays the remaining iterations for the still-live lanes only.
//
// Two instantiations of the same loop (template parameter TC):
//   TC = true  (rnnt_greedy_decode, the product path): the three contractions of an iteration — prediction-LSTM gates
//     [active lanes x 640] . [640 x 64 per CTA], g.W_pred [lanes x 320] . [320 x 26] and fc2 [16 lanes x 512] . [512 x 32]
//     — run on the tensor cores (mma.sync m16n8k16, bf16 operands from XOR-swizzled shared memory through ldmatrix, fp32
//     accumulators).  With emissions spread over the utterance ~20 % of the lanes re-evaluate the prediction network
//     per iteration (40..60 lanes per group) and the CUDA-core chains were 85 % of the kernel's time.  The fp32
//     accumulation order is the tensor core's, not the oracle's or AMX's: hg / logits agree to fp32 rounding, tokens
//     agree except at near-ties of the top-2 logits (tolerance + margin report).
//   TC = false (rnnt_greedy_decode_exact_order): the CUDA-core chains in the oracle's order described above, kept for
//     bit-exact cross-checks of the state machine and of everything downstream of a given y1.
#include "kernels.cuh"
#include "gate_math.cuh"

namespace rnnt {

enum : int { LS_CUR = 1, LS_NEED = 2, LS_FINISHED = 4, LS_FIXED = 8 };

// ---------------------------------------------------------------- mma.sync / ldmatrix (TC variant)
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// 16-byte chunk `chunk` of row `row` in a matrix of `row_bytes`-byte rows whose chunks are XOR-swizzled with the low three
// bits of the row (row_bytes % 128 == 0): the eight rows of an ldmatrix 8x8 tile then fall into eight different bank groups
__device__ __forceinline__ uint32_t swz(uint32_t base, int row, int row_bytes, int chunk) {
  return base + (uint32_t)(row * row_bytes) + (uint32_t)((chunk ^ (row & 7)) << 4);
}

__device__ __forceinline__ void group_barrier(const DecParams& p, int g, int& epoch, bool& ok, int* s_fail) {
  __syncthreads();
  ++epoch;
  if (threadIdx.x == 0 && ok) {
    int* cnt = p.group_cnt + g * 32;
    asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(cnt) : "memory");
    const int target = epoch * DEC_S;
    long long t0 = clock64(); uint32_t spins = 0;
    for (;;) {
      int v;
      asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
      if (v >= target) break;
      if ((++spins & 63u) == 0) {
        if (*(volatile int*)p.status != DEV_OK) { *s_fail = 1; break; }
        if (clock64() - t0 > kWatchdogCycles) { watchdog_fired(p.status, DEV_TIMEOUT_FLAG, 1, epoch); *s_fail = 1; break; }
      }
    }
    asm volatile("fence.acquire.gpu;" ::: "memory");
  }
  __syncthreads();
  if (*s_fail) ok = false;     // shared-memory flag: no global status read on the fast path
}

__device__ __forceinline__ void bf16x8_to_smem(float* dst, uint4 v) {
  float4 a, b;
  a.x = __uint_as_float(v.x << 16); a.y = __uint_as_float(v.x & 0xffff0000u);
  a.z = __uint_as_float(v.y << 16); a.w = __uint_as_float(v.y & 0xffff0000u);
  b.x = __uint_as_float(v.z << 16); b.y = __uint_as_float(v.z & 0xffff0000u);
  b.z = __uint_as_float(v.w << 16); b.w = __uint_as_float(v.w & 0xffff0000u);
  reinterpret_cast<float4*>(dst)[0] = a;
  reinterpret_cast<float4*>(dst)[1] = b;
}

// Dot-product chains of TWO weight rows (gate columns c and c + 32) against the activations of lanes ls, ls + 4, ... (NL
// of them; stride 4 lanes = 8 * DEC_P floats): 2 * NL independent chains per thread, every chain sequential in k from 0
// (the oracle's order).  Two columns per thread halve the activation reads per FMA — the shared-memory pipe, not the
// FMA pipe, bounded the one-column version (10.4K cycles per full chunk against 3.8K of FMA issue).
// Results to part[...][lane][col].
template <int NL>
__device__ __forceinline__ void dot_chains(const uint16_t* wrow0, const uint16_t* wrow1, const float* xbase, float* pout) {
  float acc0[NL], acc1[NL];
#pragma unroll
  for (int i = 0; i < NL; ++i) { acc0[i] = 0.f; acc1[i] = 0.f; }
  // few chains per thread = a latency-bound FMA chain: unroll deeper so that more operand loads are in flight per chain
#pragma unroll(NL <= 1 ? 8 : 4)
  for (int k4 = 0; k4 < DEC_P / 4; ++k4) {
    const uint2 wa = *reinterpret_cast<const uint2*>(wrow0 + k4 * 4);
    const uint2 wb = *reinterpret_cast<const uint2*>(wrow1 + k4 * 4);
    const float a0 = __uint_as_float(wa.x << 16), a1 = __uint_as_float(wa.x & 0xffff0000u);
    const float a2 = __uint_as_float(wa.y << 16), a3 = __uint_as_float(wa.y & 0xffff0000u);
    const float b0 = __uint_as_float(wb.x << 16), b1 = __uint_as_float(wb.x & 0xffff0000u);
    const float b2 = __uint_as_float(wb.y << 16), b3 = __uint_as_float(wb.y & 0xffff0000u);
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      const float4 xv = *reinterpret_cast<const float4*>(xbase + i * 8 * DEC_P + k4 * 4);
      acc0[i] = __fmaf_rn(xv.x, a0, acc0[i]);
      acc1[i] = __fmaf_rn(xv.x, b0, acc1[i]);
      acc0[i] = __fmaf_rn(xv.y, a1, acc0[i]);
      acc1[i] = __fmaf_rn(xv.y, b1, acc1[i]);
      acc0[i] = __fmaf_rn(xv.z, a2, acc0[i]);
      acc1[i] = __fmaf_rn(xv.z, b2, acc1[i]);
      acc0[i] = __fmaf_rn(xv.w, a3, acc0[i]);
      acc1[i] = __fmaf_rn(xv.w, b3, acc1[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < NL; ++i) { pout[4 * i * DEC_COLS] = acc0[i]; pout[4 * i * DEC_COLS + 32] = acc1[i]; }
}

// One LSTM layer for the `na` active lanes listed in `alist` (group-local lane ids), chunked by DEC_CHUNK.
//   x source: layer 0 -> embedding of pre_g (zero for SOS / out-of-table), layer 1 -> candidate h of layer 0.
//   h, c source: committed buffers (cur).  Results go to the candidate buffers (cur ^ 1).
// s_state[li] = {flags, cur_t, pre_g, fix_iter} of the group's lanes (shared-memory copy made by the list builder).
__device__ __forceinline__ void lstm_phase(const DecParams& p, int layer, int g, int s, const int* alist, int na,
                                           const int4* s_state, const uint16_t* wl, const float* s_bias, float* act,
                                           float* part, long long* lt = nullptr) {
  const int tid = threadIdx.x;
  const int lane0 = g, lstep = p.NG;      // lane li of group g is lane g + li * NG of the batch
  constexpr int VPL = 2 * DEC_P / 8;      // 16-byte vectors per lane: 40 of x, 40 of h
  for (int a0 = 0; a0 < na; a0 += DEC_CHUNK) {
    const int nc = min(DEC_CHUNK, na - a0);
    // ---- cell-update operand prefetch: thread = (unit u, lane a); its c_old is needed only after the dot products
    const int uu = tid & 15, ua = tid >> 4;
    float c_old = 0.f; int un = 0, ucur = 0;
    if (ua < nc) {
      const int li = alist[a0 + ua];
      un = lane0 + li * lstep;
      ucur = s_state[li].x & LS_CUR;
      c_old = p.cg[ucur][layer][(size_t)un * DEC_P + s * DEC_U + uu];
    }
    // ---- stage activations as fp32: act[a][0..319] = x, act[a][320..639] = h; all loads issued before any use
    {
      uint4 v[4]; int dsti[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int e = tid + r * DEC_THREADS;
        dsti[r] = -1;
        v[r] = make_uint4(0u, 0u, 0u, 0u);
        if (e < nc * VPL) {
          const int a = e / VPL, q = e % VPL;
          const int li = alist[a0 + a];
          const int4 st = s_state[li];
          const int n = lane0 + li * p.NG, cur = st.x & LS_CUR;
          dsti[r] = a * 2 * DEC_P + q * 8;
          if (q < VPL / 2) {
            if (layer == 0) {
              const int sym = st.z;
              if (sym >= 0 && sym <= 27) v[r] = __ldg(reinterpret_cast<const uint4*>(p.embed + sym * DEC_P) + q);
            } else {
              v[r] = __ldcg(reinterpret_cast<const uint4*>(p.hg[cur ^ 1][0] + (size_t)n * DEC_P) + q);
            }
          } else {
            v[r] = __ldcg(reinterpret_cast<const uint4*>(p.hg[cur][layer] + (size_t)n * DEC_P) + (q - VPL / 2));
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (dsti[r] >= 0) bf16x8_to_smem(act + dsti[r], v[r]);
    }
    __syncthreads();
    if (lt && tid == 0 && a0 == 0) lt[0] = clock64();
    // ---- dot products: thread = (column pair c / c + 32, matrix m, lane slot ls) owns the chains of lanes ls, ls + 4,
    // ls + 8; specialised by the number of lanes so that a sparse chunk issues no predicated-off instructions
    {
      const int c = tid & 31, m = (tid >> 5) & 1, ls = tid >> 6;
      const uint16_t* wrow = wl + ((size_t)(layer * 2 + m) * DEC_COLS + c) * DEC_KSTR;
      const float* xbase = act + ls * 2 * DEC_P + m * DEC_P;
      float* pout = part + (m * DEC_CHUNK + ls) * DEC_COLS + c;
      switch ((nc - ls + 3) / 4) {
        case 1: dot_chains<1>(wrow, wrow + 32 * DEC_KSTR, xbase, pout); break;
        case 2: dot_chains<2>(wrow, wrow + 32 * DEC_KSTR, xbase, pout); break;
        case 3: dot_chains<3>(wrow, wrow + 32 * DEC_KSTR, xbase, pout); break;
        default: break;
      }
    }
    __syncthreads();
    if (lt && tid == 0 && a0 == 0) lt[1] = clock64();
    // ---- cell update: thread = (unit u, lane a)
    if (ua < nc) {
      const int j = s * DEC_U + uu;
      float gate[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float gi = part[(0 * DEC_CHUNK + ua) * DEC_COLS + q * DEC_U + uu];
        const float gh = part[(1 * DEC_CHUNK + ua) * DEC_COLS + q * DEC_U + uu];
        gate[q] = __fadd_rn(__fadd_rn(gh, gi), s_bias[layer * DEC_COLS + q * DEC_U + uu]);   // float_out_accum: (acc + prev) + bias
      }
      const float it = sigmoid_ps(gate[0]), ft = sigmoid_ps(gate[1]), gt = tanh_ps(gate[2]), ot = sigmoid_ps(gate[3]);
      const float c_new = __fmaf_rn(it, gt, __fmul_rn(ft, c_old));
      const float ht = __fmul_rn(ot, tanh_ps(c_new));
      p.cg[ucur ^ 1][layer][(size_t)un * DEC_P + j] = c_new;
      p.hg[ucur ^ 1][layer][(size_t)un * DEC_P + j] = (uint16_t)float_to_bf16_bits(ht);
    }
    __syncthreads();
  }
}

// z[n][col] = sum_k g[n][k] * W_pred[col][k] for the active lanes; this CTA's columns [s*26, s*26+26) ∩ [0,512),
// whose W_pred rows live in shared memory (wp, row stride DEC_KSTR).
__device__ __forceinline__ void fc1_phase(const DecParams& p, int g, int s, const int* alist, int na, const int4* s_state,
                                          const uint16_t* wp, float* act) {
  const int tid = threadIdx.x;
  const int lane0 = g, lstep = p.NG;      // lane li of group g is lane g + li * NG of the batch
  const int col0 = s * DEC_JC, ncol = min(DEC_JC, DEC_J - col0);
  constexpr int VPL = DEC_P / 8;     // 40 vectors of 8 bf16 per lane
  for (int a0 = 0; a0 < na; a0 += 2 * DEC_CHUNK) {
    const int nc = min(2 * DEC_CHUNK, na - a0);
    {
      uint4 v[4]; int dsti[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int e = tid + r * DEC_THREADS;
        dsti[r] = -1;
        v[r] = make_uint4(0u, 0u, 0u, 0u);
        if (e < nc * VPL) {
          const int a = e / VPL, q = e % VPL;
          const int li = alist[a0 + a];
          const int cur = s_state[li].x & LS_CUR;
          dsti[r] = a * DEC_P + q * 8;
          v[r] = __ldcg(reinterpret_cast<const uint4*>(p.hg[cur ^ 1][1] + (size_t)(lane0 + li * lstep) * DEC_P) + q);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (dsti[r] >= 0) bf16x8_to_smem(act + dsti[r], v[r]);
    }
    __syncthreads();
    // thread = (column cl, lane slot lg) owns the chains of lanes lg, lg + nslot, ...: one weight read serves up to three
    // lanes (a chain per (lane, column) with its own weight read kept the shared-memory pipe busier than the FMA pipe)
    {
      const int nslot = DEC_THREADS / ncol;                   // 9 lane slots of ncol threads (19 for the last, short CTA)
      const int lg = tid / ncol, cl = tid - lg * ncol;
      if (lg < nslot && lg < nc) {
        const int nl = (nc - lg + nslot - 1) / nslot;         // <= 3 for nslot >= 8
        const uint16_t* wrow = wp + (size_t)cl * DEC_KSTR;
        const float* xr = act + lg * DEC_P;
        const int xs = nslot * DEC_P;
        float acc[3] = {0.f, 0.f, 0.f};
        if (nl == 1) {
#pragma unroll 8
          for (int k4 = 0; k4 < DEC_P / 4; ++k4) {
            const uint2 wb = *reinterpret_cast<const uint2*>(wrow + k4 * 4);
            const float4 xv = *reinterpret_cast<const float4*>(xr + k4 * 4);
            acc[0] = __fmaf_rn(xv.x, __uint_as_float(wb.x << 16), acc[0]);
            acc[0] = __fmaf_rn(xv.y, __uint_as_float(wb.x & 0xffff0000u), acc[0]);
            acc[0] = __fmaf_rn(xv.z, __uint_as_float(wb.y << 16), acc[0]);
            acc[0] = __fmaf_rn(xv.w, __uint_as_float(wb.y & 0xffff0000u), acc[0]);
          }
        } else {
#pragma unroll 4
          for (int k4 = 0; k4 < DEC_P / 4; ++k4) {
            const uint2 wb = *reinterpret_cast<const uint2*>(wrow + k4 * 4);
            const float w0 = __uint_as_float(wb.x << 16), w1 = __uint_as_float(wb.x & 0xffff0000u);
            const float w2 = __uint_as_float(wb.y << 16), w3 = __uint_as_float(wb.y & 0xffff0000u);
#pragma unroll
            for (int i = 0; i < 3; ++i) {
              if (i < nl) {
                const float4 xv = *reinterpret_cast<const float4*>(xr + i * xs + k4 * 4);
                acc[i] = __fmaf_rn(xv.x, w0, acc[i]);
                acc[i] = __fmaf_rn(xv.y, w1, acc[i]);
                acc[i] = __fmaf_rn(xv.z, w2, acc[i]);
                acc[i] = __fmaf_rn(xv.w, w3, acc[i]);
              }
            }
          }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
          if (i < nl) p.z[(size_t)(lane0 + alist[a0 + lg + i * nslot] * lstep) * DEC_J + col0 + cl] = acc[i];
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------- TC variant
// One LSTM layer for the active lanes on the tensor cores, in chunks of DEC_TC_LANES = 32 lanes (two m16 tiles):
//   act (shared, swizzled rows of 1280 B) = [lane][x 320 | h 320] bf16, copied as they are (no fp32 expansion);
//   warp w owns gate columns [8 w, 8 w + 8) of this CTA's 64: B fragments (W_ih then W_hh rows of those columns, K
//   contiguous = the "col" operand) by ldmatrix.x4 — two k-steps per instruction —, A fragments of both m tiles by
//   ldmatrix.x4, 2 x 20 k-steps x m-tiles MMAs into fp32 accumulators holding x.W_ih + h.W_hh;
//   the sums go through shared memory (overlaying act) to the cell update, thread = (unit, lane) as in the exact variant.
// x | h of one chunk of active lanes as 16-byte pieces in registers (80 per lane, 10 per thread): issued one chunk ahead so
// that their L2 round trip hides behind the previous chunk's MMAs
// Thread mapping of the staging: warp w moves the activations of the chunk's lanes 4 w .. 4 w + 3; per lane its 32 threads
// take the pieces l, l + 32 and (l < 16) l + 64 of the 80 — no index arithmetic per piece, the lane's state is one
// broadcast shared-memory read per warp.
struct LstmStage { uint4 v[4][3]; };
__device__ __forceinline__ void lstm_stage_load(const DecParams& p, int layer, int lane0, const int* alist, int a0, int nc,
                                                const int4* s_state, LstmStage& sg) {
  constexpr int HALF = DEC_P / 8;      // 40 pieces of x, then 40 of h
  const int warp = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int a = warp * 4 + i;
#pragma unroll
    for (int j = 0; j < 3; ++j) sg.v[i][j] = make_uint4(0u, 0u, 0u, 0u);
    if (a < nc) {
      const int li = alist[a0 + a];
      const int4 st = s_state[li];
      const int n = lane0 + li * p.NG, cur = st.x & LS_CUR;
      const uint4* hsrc = reinterpret_cast<const uint4*>(p.hg[cur][layer] + (size_t)n * DEC_P);
      const uint4* xsrc = nullptr;
      if (layer == 0) { if (st.z >= 0 && st.z <= 27) xsrc = reinterpret_cast<const uint4*>(p.embed + st.z * DEC_P); }
      else xsrc = reinterpret_cast<const uint4*>(p.hg[cur ^ 1][0] + (size_t)n * DEC_P);
      // piece q = l: x[l] (l < 32 < 40); q = l + 32: x[l + 32] for l < 8, else h[l - 8]; q = l + 64 (l < 16): h[l + 24]
      if (xsrc) sg.v[i][0] = (layer == 0) ? __ldg(xsrc + l) : __ldcg(xsrc + l);
      if (l < 8) { if (xsrc) sg.v[i][1] = (layer == 0) ? __ldg(xsrc + l + 32) : __ldcg(xsrc + l + 32); }
      else sg.v[i][1] = __ldcg(hsrc + (l - 8));
      if (l < 16) sg.v[i][2] = __ldcg(hsrc + (l + 24));
    }
  }
}
__device__ __forceinline__ void lstm_stage_store(uint8_t* act_g, int nc, const LstmStage& sg) {
  const int warp = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int a = warp * 4 + i;
    if (a < nc) {
      uint8_t* row = act_g + a * (2 * DEC_P * 2);
      const int x7 = a & 7;
      *reinterpret_cast<uint4*>(row + ((l ^ x7) << 4)) = sg.v[i][0];
      *reinterpret_cast<uint4*>(row + (((l + 32) ^ x7) << 4)) = sg.v[i][1];
      if (l < 16) *reinterpret_cast<uint4*>(row + (((l + 64) ^ x7) << 4)) = sg.v[i][2];
    }
  }
}

// The contraction of one chunk of lanes on the tensor cores.  Warp w = (matrix m = w >> 2: x.W_ih or h.W_hh, gate-column
// tiles 2 (w & 3) and 2 (w & 3) + 1): an A fragment (ldmatrix.x4 of the lanes' activations) serves two n tiles, a B
// fragment two m tiles.  ldmatrix is a volatile asm (it must stay behind the barrier that publishes the staged rows), so
// the order below is the order of the SASS: the fragments of k-pair kp + 1 are requested BEFORE the MMAs of k-pair kp are
// issued (two fragment sets in registers) — issued in program order as load, MMA, MMA, load ... every ldmatrix latency sat
// on the critical path (4.7 K cycles per chunk for 140 instructions, profiles/r02/dec_trace_tc_runahead.txt).  MT = m16
// tiles with live lanes; rows of a tile past the chunk's lanes hold stale bytes, their results are never read.
template <int MT>
__device__ __forceinline__ void lstm_dot_tc(uint32_t wl_s, uint32_t act_s, int layer, float (&acc)[2][2][2][4]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m = warp >> 2, np = warp & 3;
  const int brow = (layer * 2 + m) * DEC_COLS + 2 * np * 8 + (lane & 7);
  const int arow = (lane & 7) + ((lane >> 3) & 1) * 8;
  uint32_t b[2][2][4], a[2][2][MT][4];
  auto load = [&](int kp, int buf) {
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) ldsm_x4(swz(wl_s, brow + nt * 8, DEC_P * 2, kp * 4 + (lane >> 3)), b[buf][nt]);
#pragma unroll
    for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
      for (int mi = 0; mi < MT; ++mi)
        ldsm_x4(swz(act_s, mi * 16 + arow, 2 * DEC_P * 2, m * (DEC_P / 8) + (kp * 2 + k2) * 2 + (lane >> 4)), a[buf][k2][mi]);
  };
  load(0, 0);
#pragma unroll
  for (int kp = 0; kp < DEC_P / 32; ++kp) {
    const int buf = kp & 1;
    if (kp + 1 < DEC_P / 32) load(kp + 1, buf ^ 1);
#pragma unroll
    for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
      for (int mi = 0; mi < MT; ++mi)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) mma_bf16_16816(acc[nt][mi][k2], a[buf][k2][mi], b[buf][nt][2 * k2], b[buf][nt][2 * k2 + 1]);
  }
}

__device__ __forceinline__ void lstm_phase_tc(const DecParams& p, int layer, int g, int s, const int* alist, int na,
                                              const int4* s_state, uint32_t wl_s, const float* s_bias, uint8_t* act_g,
                                              uint32_t act_s, long long* lt = nullptr) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int lane0 = g, lstep = p.NG;      // lane li of group g is lane g + li * NG of the batch
  float* part = reinterpret_cast<float*>(act_g);          // [ih | hh][32 lanes][64 columns] fp32 partial sums, after the MMAs
  LstmStage sg;
  lstm_stage_load(p, layer, lane0, alist, 0, min(DEC_TC_LANES, na), s_state, sg);
  for (int a0 = 0; a0 < na; a0 += DEC_TC_LANES) {
    const int nc = min(DEC_TC_LANES, na - a0);
    const int mt = (nc + 15) >> 4;
    lstm_stage_store(act_g, nc, sg);
    __syncthreads();
    if (a0 + DEC_TC_LANES < na) lstm_stage_load(p, layer, lane0, alist, a0 + DEC_TC_LANES, min(DEC_TC_LANES, na - a0 - DEC_TC_LANES), s_state, sg);
    // ---- cell-update operand prefetch: thread = (unit uu, lanes ua and ua + 16)
    const int uu = tid & 15, ua = tid >> 4;
    float c_old[2] = {0.f, 0.f}; int un[2] = {0, 0}, ucur[2] = {0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int a = ua + 16 * h;
      if (a < nc) {
        const int li = alist[a0 + a];
        un[h] = lane0 + li * lstep;
        ucur[h] = s_state[li].x & LS_CUR;
        c_old[h] = p.cg[ucur[h]][layer][(size_t)un[h] * DEC_P + s * DEC_U + uu];
      }
    }
    if (lt && tid == 0 && a0 == 0) lt[0] = clock64();
    // ---- tensor-core contraction (lstm_dot_tc); accumulators per (n tile, m tile, k-step parity) are independent chains
    float acc[2][2][2][4];
#pragma unroll
    for (int i = 0; i < 32; ++i) (&acc[0][0][0][0])[i] = 0.f;
    if (mt == 2) lstm_dot_tc<2>(wl_s, act_s, layer, acc);
    else lstm_dot_tc<1>(wl_s, act_s, layer, acc);
    __syncthreads();                       // every warp is done reading act: reuse it for the partial gate sums
    {                                       // part[m][lane][col] fp32
      const int m = warp >> 2, np = warp & 3;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
          if (mi < mt) {
            const int r0 = mi * 16 + (lane >> 2), c0 = (2 * np + nt) * 8 + 2 * (lane & 3);
            float* dst = part + (m * DEC_TC_LANES + r0) * DEC_COLS + c0;
            *reinterpret_cast<float2*>(dst) = make_float2(acc[nt][mi][0][0] + acc[nt][mi][1][0], acc[nt][mi][0][1] + acc[nt][mi][1][1]);
            *reinterpret_cast<float2*>(dst + 8 * DEC_COLS) = make_float2(acc[nt][mi][0][2] + acc[nt][mi][1][2], acc[nt][mi][0][3] + acc[nt][mi][1][3]);
          }
        }
    }
    __syncthreads();
    if (lt && tid == 0 && a0 == 0) lt[1] = clock64();
    // ---- cell update: thread = (unit uu, lanes ua, ua + 16)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int a = ua + 16 * h;
      if (a < nc) {
        const int j = s * DEC_U + uu;
        float gate[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)        // float_out_accum: (h.W_hh + x.W_ih) + bias
          gate[q] = __fadd_rn(__fadd_rn(part[(DEC_TC_LANES + a) * DEC_COLS + q * DEC_U + uu], part[a * DEC_COLS + q * DEC_U + uu]),
                              s_bias[layer * DEC_COLS + q * DEC_U + uu]);
        const float it = sigmoid_ps(gate[0]), ft = sigmoid_ps(gate[1]), gt = tanh_ps(gate[2]), ot = sigmoid_ps(gate[3]);
        const float c_new = __fmaf_rn(it, gt, __fmul_rn(ft, c_old[h]));
        const float ht = __fmul_rn(ot, tanh_ps(c_new));
        p.cg[ucur[h] ^ 1][layer][(size_t)un[h] * DEC_P + j] = c_new;
        p.hg[ucur[h] ^ 1][layer][(size_t)un[h] * DEC_P + j] = (uint16_t)float_to_bf16_bits(ht);
      }
    }
    __syncthreads();
  }
}

// z[n][col] = sum_k g[n][k] * W_pred[col][k] on the tensor cores: chunk of 32 lanes x this CTA's <= 26 columns (four n8
// tiles; the rows past ncol of the B tile alias row ncol - 1 and their results are dropped).
__device__ __forceinline__ void fc1_phase_tc(const DecParams& p, int g, int s, const int* alist, int na, const int4* s_state,
                                             uint32_t wp_s, uint8_t* act_g, uint32_t act_s) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int lane0 = g, lstep = p.NG;      // lane li of group g is lane g + li * NG of the batch
  const int col0 = s * DEC_JC, ncol = min(DEC_JC, DEC_J - col0);
  // g = h1 of a chunk's lanes as 16-byte pieces in registers, requested one chunk ahead (their L2 round trip hides behind
  // the previous chunk's MMAs, as in lstm_phase_tc)
  constexpr int VPL = DEC_P / 8;          // 40
  constexpr int R = DEC_TC_LANES * VPL / DEC_THREADS;   // 5
  uint4 v[R]; int dst[R];
  auto stage_load = [&](int a0, int nc) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int e = tid + r * DEC_THREADS;
      const int a = e / VPL, q = e % VPL;
      dst[r] = -1;
      v[r] = make_uint4(0u, 0u, 0u, 0u);
      if (a < nc) {
        const int li = alist[a0 + a];
        const int cur = s_state[li].x & LS_CUR;
        dst[r] = a * (2 * DEC_P * 2) + ((q ^ (a & 7)) << 4);
        v[r] = __ldcg(reinterpret_cast<const uint4*>(p.hg[cur ^ 1][1] + (size_t)(lane0 + li * lstep) * DEC_P) + q);
      }
    }
  };
  if (na > 0) stage_load(0, min(DEC_TC_LANES, na));
  for (int a0 = 0; a0 < na; a0 += DEC_TC_LANES) {
    const int nc = min(DEC_TC_LANES, na - a0);
    const int mt = (nc + 15) >> 4;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (dst[r] >= 0) *reinterpret_cast<uint4*>(act_g + dst[r]) = v[r];
    __syncthreads();
    if (a0 + DEC_TC_LANES < na) stage_load(a0 + DEC_TC_LANES, min(DEC_TC_LANES, na - a0 - DEC_TC_LANES));
    const int mi = warp >> 2, ni = warp & 3;
    if (mi < mt) {
      // fragments of k-pair kp + 1 requested before the MMAs of k-pair kp (see lstm_dot_tc); one accumulator per k-step
      // parity, summed at the end
      float acc2[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      const int brow = min(ni * 8 + (lane & 7), ncol - 1);
      const int arow = mi * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
      uint32_t b[2][4], a[2][2][4];
      auto load = [&](int kp, int buf) {
        ldsm_x4(swz(wp_s, brow, DEC_P * 2, kp * 4 + (lane >> 3)), b[buf]);
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) ldsm_x4(swz(act_s, arow, 2 * DEC_P * 2, (kp * 2 + k2) * 2 + (lane >> 4)), a[buf][k2]);
      };
      load(0, 0);
#pragma unroll
      for (int kp = 0; kp < DEC_P / 32; ++kp) {
        const int buf = kp & 1;
        if (kp + 1 < DEC_P / 32) load(kp + 1, buf ^ 1);
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) mma_bf16_16816(acc2[k2], a[buf][k2], b[buf][2 * k2], b[buf][2 * k2 + 1]);
      }
      float acc[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = acc2[0][i] + acc2[1][i];
      const int r0 = mi * 16 + (lane >> 2), c0 = ni * 8 + 2 * (lane & 3);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = r0 + 8 * h;
        if (r < nc) {
          float* zr = p.z + (size_t)(lane0 + alist[a0 + r] * lstep) * DEC_J + col0;
          if (c0 < ncol) zr[c0] = acc[2 * h];
          if (c0 + 1 < ncol) zr[c0 + 1] = acc[2 * h + 1];
        }
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------- phase D (joint + update)
// A lane's decisions depend on nothing but its own state, so phase D lets every lane RUN AHEAD: as long as its
// decisions are blank (or the forced advance after 30 symbols) it only needs the next transcription row y1[t] — z, the
// prediction half of fc1, is unchanged — and takes decision after decision without any group barrier.  A lane stops
// when it accepts a symbol (the prediction network must be re-evaluated: next round), when it finishes, or — in the
// replay phase of quirk (b) — when it has taken as many decisions as the batch-lockstep loop of the reference runs.
// The per-lane decision count `dc` (lane_state.w) is what the reference's loop index would be.
struct JointRegs { int sa, ri, ti, flen; };

// y = bf16(relu((z + y1[t]) + b1)) for k = 16 lane .. 16 lane + 15 of one lane -> 32 bytes of its row in shared memory.
__device__ __forceinline__ void joint_y(const float (&zf)[16], const uint4 (&yv)[2], const float (&bf)[16], uint16_t* yrow, int lane) {
  const uint32_t* yw = reinterpret_cast<const uint32_t*>(yv);
  uint32_t out[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float y0 = __uint_as_float(yw[i] << 16), y1f = __uint_as_float(yw[i] & 0xffff0000u);
    const float v0 = __fadd_rn(__fadd_rn(zf[2 * i], y0), bf[2 * i]);            // (acc + y1) + bias
    const float v1 = __fadd_rn(__fadd_rn(zf[2 * i + 1], y1f), bf[2 * i + 1]);
    out[i] = float_to_bf16_bits(fmaxf(v0, 0.f)) | (float_to_bf16_bits(fmaxf(v1, 0.f)) << 16);   // relu, round to bf16
  }
  uint4* yo = reinterpret_cast<uint4*>(yrow + lane * 16);
  yo[0] = make_uint4(out[0], out[1], out[2], out[3]);
  yo[1] = make_uint4(out[4], out[5], out[6], out[7]);
}
__device__ __forceinline__ void load_y1(const DecParams& p, int n, int t, uint4 (&yv)[2], int lane) {
  const uint4* src = reinterpret_cast<const uint4*>(p.y1 + ((size_t)t * p.N + n) * DEC_J + lane * 16);
  yv[0] = __ldg(src);
  yv[1] = __ldg(src + 1);
}

// argmax over the 32 logits (lane = column, first maximum wins ties like torch.argmax) + the greedy_decode_update state
// machine of one decision (greedy_decode_update.cpp:13-96).  `st` / `r` are updated in place (all threads of the warp
// hold the same st afterwards); returns true if the lane may take another decision right away.
__device__ __forceinline__ bool joint_update(const DecParams& p, int n, int4& st, JointRegs& r, float logit, int lane, int total) {
  float best = logit;
  int bidx = lane;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {   // first maximum wins ties (torch.argmax)
    const float ov = __shfl_xor_sync(0xffffffffu, best, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
    if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
  }
  const int sym = bidx;                       // identical in every lane of the warp
  int flags = st.x, cur_t = st.y, pre_g = st.z;
  const int dc = st.w + 1;
  const bool upd_g = (sym != DEC_BLANK) && (r.sa != DEC_MAXSYM) && (r.flen != 0);
  if (upd_g) {
    r.ri += 1;
    r.sa += 1;
    if (lane == 0 && r.ri >= 0 && r.ri < p.L) p.res[(size_t)n * p.L + r.ri] = sym;
    pre_g = sym;
    flags ^= LS_CUR;       // candidate becomes the committed state
    flags |= LS_NEED;      // prediction network must be re-evaluated
  } else {
    r.sa = 0;
    r.ti += 1;
    flags &= ~LS_NEED;
  }
  const bool finished = r.ti >= r.flen;
  if (!finished) cur_t = r.ti;                // fi = f[time_idx]
  flags = finished ? (flags | LS_FINISHED) : (flags & ~LS_FINISHED);
  if (finished && !upd_g && (sym == DEC_BLANK || r.flen == 0)) flags |= LS_FIXED;   // fixed point of the reference loop
  st = make_int4(flags, cur_t, pre_g, dc);
  if (upd_g || (flags & LS_FIXED)) return false;
  if (finished) return total >= 0 && dc < total;        // capped on its last frame: parked until the replay phase
  return true;
}
// r.* live in every thread of the warp (loaded by lane 0 and broadcast); lane 0 writes them back
__device__ __forceinline__ JointRegs load_counters(const DecParams& p, int n, int lane) {
  JointRegs r = {0, 0, 0, 0};
  if (lane == 0) { r.sa = p.symbols_added[n]; r.ri = p.res_idx[n]; r.ti = p.time_idx[n]; r.flen = p.f_lens[n]; }
  r.sa = __shfl_sync(0xffffffffu, r.sa, 0); r.ri = __shfl_sync(0xffffffffu, r.ri, 0);
  r.ti = __shfl_sync(0xffffffffu, r.ti, 0); r.flen = __shfl_sync(0xffffffffu, r.flen, 0);
  return r;
}
// `next` = the lane-state copy the NEXT round's scan reads (see the round loop: the states are double-buffered)
__device__ __forceinline__ void store_lane(const DecParams& p, int32_t* next, int n, int4 st, JointRegs r, int lane) {
  if (lane == 0) {
    p.symbols_added[n] = r.sa; p.res_idx[n] = r.ri; p.time_idx[n] = r.ti; p.pre_g[n] = st.z;
    *reinterpret_cast<int4*>(next + (size_t)n * 4) = st;
  }
}
// may this lane take decisions in this round?  phase 1 (total < 0): every unfinished lane; replay phase: the lanes that
// finished on the 30-symbol cap, until they have taken `total` decisions
__device__ __forceinline__ bool runnable(int4 st, int total) {
  if (st.x & LS_FIXED) return false;
  if (total < 0) return !(st.x & LS_FINISHED);
  return (st.x & LS_FINISHED) && st.w < total;
}

// exact variant: logits[lane] = y . W2[lane]^T as one sequential chain over k (the oracle's order), W2 rows in shared memory
__device__ __forceinline__ float fc2_chain(const uint8_t* w2s, const uint16_t* ybuf, int lane) {
  float acc = 0.f;
  if (lane < DecSmem::W2_ROWS) {
    const uint4* wrow = reinterpret_cast<const uint4*>(w2s + lane * DecSmem::W2_STRIDE);
    const uint4* yrow = reinterpret_cast<const uint4*>(ybuf);
#pragma unroll 4
    for (int k8 = 0; k8 < DEC_J / 8; ++k8) {
      const uint4 wb = wrow[k8];
      const uint4 yv = yrow[k8];
      acc = __fmaf_rn(__uint_as_float(yv.x << 16), __uint_as_float(wb.x << 16), acc);
      acc = __fmaf_rn(__uint_as_float(yv.x & 0xffff0000u), __uint_as_float(wb.x & 0xffff0000u), acc);
      acc = __fmaf_rn(__uint_as_float(yv.y << 16), __uint_as_float(wb.y << 16), acc);
      acc = __fmaf_rn(__uint_as_float(yv.y & 0xffff0000u), __uint_as_float(wb.y & 0xffff0000u), acc);
      acc = __fmaf_rn(__uint_as_float(yv.z << 16), __uint_as_float(wb.z << 16), acc);
      acc = __fmaf_rn(__uint_as_float(yv.z & 0xffff0000u), __uint_as_float(wb.z & 0xffff0000u), acc);
      acc = __fmaf_rn(__uint_as_float(yv.w << 16), __uint_as_float(wb.w << 16), acc);
      acc = __fmaf_rn(__uint_as_float(yv.w & 0xffff0000u), __uint_as_float(wb.w & 0xffff0000u), acc);
    }
  }
  return acc;
}

template <bool TC>
__global__ void __launch_bounds__(DEC_THREADS, 1) greedy_decode_kernel(const DecParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  // exact variant: wl | act | part | pad | wp (padded rows, fp32 activations); TC variant: wl | act | wp (swizzled bf16)
  constexpr int W_BYTES = TC ? DecSmemTC::W_BYTES : DecSmem::W_BYTES;
  constexpr int WP_OFFSET = TC ? DecSmemTC::W_BYTES + DecSmemTC::ACT_BYTES
                               : DecSmem::W_BYTES + DecSmem::ACT_BYTES + DecSmem::PART_BYTES + DecSmem::PAD_BYTES;
  uint16_t* wl = reinterpret_cast<uint16_t*>(smem);
  float* act = reinterpret_cast<float*>(smem + W_BYTES);
  float* part = reinterpret_cast<float*>(smem + DecSmem::W_BYTES + DecSmem::ACT_BYTES);       // exact variant only
  uint16_t* wp = reinterpret_cast<uint16_t*>(smem + WP_OFFSET);
  uint8_t* w2s = smem + W_BYTES;                                             // exact: joint-phase overlay of act | part | pad
  uint16_t* jy = reinterpret_cast<uint16_t*>(w2s + DecSmem::JY_OFFSET);
  const uint32_t wl_s = smem_u32(smem), act_s = smem_u32(smem + W_BYTES), wp_s = smem_u32(smem + WP_OFFSET);
  __shared__ int4 s_state[DEC_MAXLG];
  __shared__ int alist[DEC_MAXLG];
  __shared__ float s_bias[2 * DEC_COLS], s_b2[DEC_V];
  __shared__ int4 s_blk[(DEC_MAXLG + 31) / 32];      // per 32-lane block: need mask, unfinished, live, max decisions
  __shared__ int s_unfinished, s_fail;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = blockIdx.x / DEC_S, s = blockIdx.x % DEC_S;
  const int lane0 = g, lstep = p.NG;      // lane li of group g is lane g + li * NG of the batch
  const int nl = g < p.N ? (p.N - g + p.NG - 1) / p.NG : 0;    // lanes of this group
  const int col0 = s * DEC_JC, ncol = min(DEC_JC, DEC_J - col0);

  // resident weights: wl[layer][m][c][k], c = gate*16 + u  <-  W[(gate*320 + s*16 + u)][k]
  if constexpr (TC) {
    constexpr int CPR = DEC_P / 8;        // 40 16-byte chunks per row
    for (int e = tid; e < 2 * 2 * DEC_COLS * CPR; e += DEC_THREADS) {
      const int q = e % CPR, r = e / CPR;                 // r = (layer * 2 + m) * 64 + c
      const int c = r % DEC_COLS, m = (r / DEC_COLS) & 1, layer = r / (2 * DEC_COLS);
      const uint16_t* src = (m == 0 ? p.w_ih[layer] : p.w_hh[layer]) + (size_t)((c / DEC_U) * DEC_P + s * DEC_U + (c % DEC_U)) * DEC_P;
      *reinterpret_cast<uint4*>(smem + r * (DEC_P * 2) + ((q ^ (r & 7)) << 4)) = __ldg(reinterpret_cast<const uint4*>(src) + q);
    }
    for (int e = tid; e < ncol * CPR; e += DEC_THREADS) {
      const int q = e % CPR, cl = e / CPR;
      *reinterpret_cast<uint4*>(smem + WP_OFFSET + cl * (DEC_P * 2) + ((q ^ (cl & 7)) << 4)) =
          __ldg(reinterpret_cast<const uint4*>(p.w_pred + (size_t)(col0 + cl) * DEC_P) + q);
    }
  } else {
    for (int e = tid; e < 2 * 2 * DEC_COLS * (DEC_P / 2); e += DEC_THREADS) {
      const int k2 = e % (DEC_P / 2);
      int r = e / (DEC_P / 2);
      const int c = r % DEC_COLS; r /= DEC_COLS;
      const int m = r & 1, layer = r >> 1;
      const uint16_t* src = (m == 0 ? p.w_ih[layer] : p.w_hh[layer]) + (size_t)((c / DEC_U) * DEC_P + s * DEC_U + (c % DEC_U)) * DEC_P;
      reinterpret_cast<uint32_t*>(wl + ((size_t)(layer * 2 + m) * DEC_COLS + c) * DEC_KSTR)[k2] =
          reinterpret_cast<const uint32_t*>(src)[k2];
    }
    for (int e = tid; e < ncol * (DEC_P / 2); e += DEC_THREADS) {
      const int k2 = e % (DEC_P / 2), cl = e / (DEC_P / 2);
      reinterpret_cast<uint32_t*>(wp + (size_t)cl * DEC_KSTR)[k2] =
          reinterpret_cast<const uint32_t*>(p.w_pred + (size_t)(col0 + cl) * DEC_P)[k2];
    }
  }
  for (int e = tid; e < 2 * DEC_COLS; e += DEC_THREADS) {
    const int layer = e / DEC_COLS, c = e % DEC_COLS;
    s_bias[e] = p.bias[layer][(c / DEC_U) * DEC_P + s * DEC_U + (c % DEC_U)];
  }
  if (tid < DEC_V) s_b2[tid] = p.b2[tid];
  // fc1 bias of this thread's 16 columns (phase D: thread = columns 16 lane .. 16 lane + 15 of its warp's lanes)
  float b1r[16];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p.b1 + lane * 16) + i);
    b1r[4 * i] = v.x; b1r[4 * i + 1] = v.y; b1r[4 * i + 2] = v.z; b1r[4 * i + 3] = v.w;
  }
  // TC variant: fc2 B fragments stay in registers for the whole kernel: warp = (n tile w & 3, K half w >> 2), 16 k-steps
  uint32_t w2f[TC ? 32 : 1];
  if constexpr (TC) {
    const uint16_t* wrow = p.w2 + (size_t)((warp & 3) * 8 + (lane >> 2)) * DEC_J + (warp >> 2) * (DEC_J / 2) + 2 * (lane & 3);
#pragma unroll
    for (int ks = 0; ks < 16; ++ks) {
      w2f[2 * ks] = __ldg(reinterpret_cast<const uint32_t*>(wrow + ks * 16));
      w2f[2 * ks + 1] = __ldg(reinterpret_cast<const uint32_t*>(wrow + ks * 16 + 8));
    }
  }
  // lane state init (each lane by the CTA that owns it in phase D)
  for (int li = s + tid * DEC_S; li < nl; li += DEC_S * DEC_THREADS)
    *reinterpret_cast<int4*>(p.lane_state + (size_t)(lane0 + li * lstep) * 4) = make_int4(LS_NEED, 0, p.pre_g[lane0 + li * lstep], 0);
  if (tid == 0) s_fail = 0;
  int epoch = 0; bool ok = true;
  group_barrier(p, g, epoch, ok, &s_fail);

  int round = 0;
  // Lane states are double-buffered by round parity: the scan of a round reads copy `par`, the decisions of the round write
  // copy `par ^ 1` (every CTA carries the states of its lanes that took no decision forward).  A round without prediction
  // work has no group barrier between the scan and the decisions, so with a single copy a CTA that is a few microseconds
  // late in its scan could read a state another CTA has already advanced, build a different work list and fall out of
  // step with its group; with two copies the scan of round r never sees a store of round r (and a CTA cannot be two rounds
  // ahead: one barrier per round).
  int par = 0;
  int total = -1;         // batch-lockstep iteration count of the reference loop; unknown during phase 1
  bool w2_valid = false;
  for (;;) {
    // ---- snapshot the group's lane states (one 16-byte load per lane), build the list of lanes whose prediction
    // network has to be re-evaluated, count the lanes still to finish (phase 1) / still to replay (phase 2).  This sits on
    // the critical path of every round (a lane that emits symbol after symbol is a chain of rounds), so it is two
    // barriers deep: warp w scans the 32-lane blocks w, w + 8 straight from global memory, the per-block results go
    // through shared memory, every thread sums the <= 10 of them.
#ifdef RNNT_DEBUG_HOOKS
    if (p.skew > 0 && s == 1) {       // one CTA of the group scans late: the other CTAs' decisions of this round are under way
      const long long t0 = clock64();
      while (clock64() - t0 < p.skew) {}
    }
#endif
    const int nblk = (nl + 31) >> 5;
    bool need_r[2] = {false, false};
    for (int b = warp, i = 0; b < nblk; b += DEC_WARPS, ++i) {
      const int li = b * 32 + lane;
      int need = 0, unfin = 0, live = 0, dc = 0;
      if (li < nl) {
        const int4 st = __ldcg(reinterpret_cast<const int4*>(p.lane_state + ((size_t)par * p.N + lane0 + li * lstep) * 4));
        s_state[li] = st;
        const bool replay = total >= 0 && (st.x & LS_FINISHED) && !(st.x & LS_FIXED) && st.w < total;
        unfin = !(st.x & LS_FINISHED);
        live = unfin || replay;
        need = (st.x & LS_NEED) && !(st.x & LS_FIXED) && live;
        dc = st.w;
      }
      const unsigned mneed = __ballot_sync(0xffffffffu, need);
      const unsigned munf = __ballot_sync(0xffffffffu, unfin);
      const unsigned mlive = __ballot_sync(0xffffffffu, live);
      dc = __reduce_max_sync(0xffffffffu, dc);
      need_r[i] = need != 0;
      if (lane == 0) s_blk[b] = make_int4((int)mneed, __popc(munf), __popc(mlive), dc);
    }
    __syncthreads();
    int na = 0, unfinished = 0, live_lanes = 0, maxdc = 0, before[2] = {0, 0};
    for (int b = 0; b < nblk; ++b) {
      const int4 v = s_blk[b];
      if (b == warp) before[0] = na;
      if (b == warp + DEC_WARPS) before[1] = na;
      na += __popc((unsigned)v.x); unfinished += v.y; live_lanes += v.z; maxdc = max(maxdc, v.w);
    }
    for (int b = warp, i = 0; b < nblk; b += DEC_WARPS, ++i)
      if (need_r[i]) alist[before[i] + __popc((unsigned)s_blk[b].x & ((1u << lane) - 1u))] = b * 32 + lane;
    __syncthreads();
    if (!ok) break;
    if (total < 0 && unfinished == 0) {
      // ---- every lane of the group has finished: rendezvous of all groups; the largest decision count any lane needed to
      // finish is the number of iterations of the reference's lock-step loop
      if (tid == 0) {
        if (s == 0) {
          atomicMax(p.global_sync + 1, maxdc);
          __threadfence();
          atomicAdd(p.global_sync + 0, 1);
        }
        long long t0 = clock64(); uint32_t spins = 0;
        for (;;) {
          int v;
          asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p.global_sync) : "memory");
          if (v >= p.NG) break;
          if ((++spins & 63u) == 0) {
            if (*(volatile int*)p.status != DEV_OK) break;
            if (clock64() - t0 > kWatchdogCycles) { watchdog_fired(p.status, DEV_TIMEOUT_FLAG, 2, g); break; }
          }
        }
        s_unfinished = *(volatile int*)(p.global_sync + 1);
      }
      __syncthreads();
      total = s_unfinished;
      __syncthreads();
      if (*(volatile int*)p.status != DEV_OK) { ok = false; break; }
      continue;            // rebuild the lists with `total` known: the lanes parked on the cap become live again
    }
    if (total >= 0 && live_lanes == 0) break;
    ++round;
    long long* const trace = RNNT_TRACE_PTR(p);
    const bool tr = trace && (int)blockIdx.x == p.trace_block && tid == 0 && round <= p.trace_cap;
    long long* trow = trace + (size_t)(round - 1) * 16;
    if (tr) { trow[0] = clock64(); trow[7] = na; }

    if (na > 0) {
      if constexpr (TC) lstm_phase_tc(p, 0, g, s, alist, na, s_state, wl_s, s_bias, smem + W_BYTES, act_s, tr ? trow + 12 : nullptr);
      else lstm_phase(p, 0, g, s, alist, na, s_state, wl, s_bias, act, part, tr ? trow + 12 : nullptr);
      if (tr) trow[1] = clock64();
      group_barrier(p, g, epoch, ok, &s_fail);
      if (tr) trow[2] = clock64();
      if constexpr (TC) lstm_phase_tc(p, 1, g, s, alist, na, s_state, wl_s, s_bias, smem + W_BYTES, act_s);
      else lstm_phase(p, 1, g, s, alist, na, s_state, wl, s_bias, act, part);
      group_barrier(p, g, epoch, ok, &s_fail);
      if (tr) trow[3] = clock64();
      if constexpr (TC) fc1_phase_tc(p, g, s, alist, na, s_state, wp_s, smem + W_BYTES, act_s);
      else fc1_phase(p, g, s, alist, na, s_state, wp, act);
      group_barrier(p, g, epoch, ok, &s_fail);
      if (tr) trow[4] = clock64();
    }
    int decisions = 0;
    int32_t* const ls_next = p.lane_state + (size_t)(par ^ 1) * p.N * 4;
    if constexpr (TC) {
      // ---- phase D on the tensor cores: this CTA's lanes li = s + 20 j (j < 16) form ONE m16 tile; warp w drives the
      // lanes j = w and w + 8.  Sub-iteration: every running lane stages y of its current frame as a row of a [16][512]
      // bf16 matrix (row pitch 1040 B: ldmatrix rows in different bank groups); all 8 warps multiply it with W2
      // (warp = n tile x K half, B fragments in registers); the two partial sums + b2 go through shared memory to the
      // lanes' warps, which take the decision; repeat while any lane of the CTA keeps running.
      uint8_t* jy_g = smem + W_BYTES;
      float* lg = reinterpret_cast<float*>(smem + W_BYTES + DecSmemTC::JY_BYTES);       // [K half][16][32] fp32
      JointRegs jr[2]; int4 st2[2]; bool run[2]; float zf[2][16]; uint4 yv[2][2], yn[2][2];
      int budget = p.max_run;          // decisions a lane may take per round: keeps the CTAs of a group in step
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = warp + 8 * h, li = s + j * DEC_S;
        run[h] = false;
        jr[h] = JointRegs{0, 0, 0, 0};
        st2[h] = make_int4(LS_FIXED, 0, 0, 0);
        if (li < nl) {
          st2[h] = s_state[li];
          run[h] = runnable(st2[h], total);
          if (!run[h] && lane == 0) *reinterpret_cast<int4*>(ls_next + (size_t)(lane0 + li * lstep) * 4) = st2[h];      // carried forward
        }
        if (run[h]) {
          const int n = lane0 + li * lstep;
          jr[h] = load_counters(p, n, lane);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 v = __ldcg(reinterpret_cast<const float4*>(p.z + (size_t)n * DEC_J + lane * 16) + i);
            zf[h][4 * i] = v.x; zf[h][4 * i + 1] = v.y; zf[h][4 * i + 2] = v.z; zf[h][4 * i + 3] = v.w;
          }
          load_y1(p, n, st2[h].y, yv[h], lane);
          load_y1(p, n, min(st2[h].y + 1, p.T2 - 1), yn[h], lane);      // the frame after: needed if this decision is a blank
        }
      }
      const uint32_t arow = act_s + (uint32_t)(((lane & 7) + ((lane >> 3) & 1) * 8) * DecSmemTC::JY_PITCH) +
                            (uint32_t)((warp >> 2) * DEC_J) + (uint32_t)((lane >> 4) << 4);
      bool first_it = true;
      while (__syncthreads_or(run[0] || run[1])) {
        --budget;
        if (tr && first_it) trow[9] = clock64();
#pragma unroll
        for (int h = 0; h < 2; ++h)
          if (run[h]) joint_y(zf[h], yv[h], b1r, reinterpret_cast<uint16_t*>(jy_g + (warp + 8 * h) * DecSmemTC::JY_PITCH), lane);
        __syncthreads();
        if (tr && first_it) trow[10] = clock64();
        {
          float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {       // four fragments in flight per batch of MMAs (ldmatrix is a volatile asm)
            uint32_t a[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i) ldsm_x4(arow + (k4 * 4 + i) * 32, a[i]);
#pragma unroll
            for (int i = 0; i < 4; ++i) mma_bf16_16816(acc[i & 1], a[i], w2f[2 * (k4 * 4 + i)], w2f[2 * (k4 * 4 + i) + 1]);
          }
          float* out = lg + (warp >> 2) * (16 * DEC_V);
          const int r0 = lane >> 2, c0 = (warp & 3) * 8 + 2 * (lane & 3);
          *reinterpret_cast<float2*>(out + r0 * DEC_V + c0) = make_float2(acc[0][0] + acc[1][0], acc[0][1] + acc[1][1]);
          *reinterpret_cast<float2*>(out + (r0 + 8) * DEC_V + c0) = make_float2(acc[0][2] + acc[1][2], acc[0][3] + acc[1][3]);
        }
        __syncthreads();
        if (tr && first_it) trow[11] = clock64();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (run[h]) {
            const int j = warp + 8 * h, n = lane0 + (s + j * DEC_S) * lstep;
            const float logit = __fadd_rn(__fadd_rn(lg[j * DEC_V + lane], lg[16 * DEC_V + j * DEC_V + lane]), s_b2[lane]);
            const int t_before = st2[h].y;
            run[h] = joint_update(p, n, st2[h], jr[h], logit, lane, total) && budget > 0;
            ++decisions;
            if (run[h] && st2[h].y != t_before) {        // advanced by one frame: the prefetched row becomes current
              yv[h][0] = yn[h][0]; yv[h][1] = yn[h][1];
              load_y1(p, n, min(st2[h].y + 1, p.T2 - 1), yn[h], lane);
            }
            if (!run[h]) store_lane(p, ls_next, n, st2[h], jr[h], lane);
          }
        }
        if (tr && first_it) trow[14] = clock64();
        first_it = false;
      }
      if (tr) trow[15] = clock64();
    } else {
      // ---- phase D, exact variant: lanes li = s, s + S, ...; one warp per lane, decision after decision.  The LSTM / fc1
      // phases overwrite the overlay, so the fc2 weights are staged again whenever they ran.
      if (na > 0) w2_valid = false;
      if (!w2_valid) {
        constexpr int NV = DecSmem::W2_ROWS * (DEC_J / 8);       // 16-byte vectors of W2 rows 0..28
        for (int e = tid; e < NV; e += DEC_THREADS)
          *reinterpret_cast<uint4*>(w2s + (e / (DEC_J / 8)) * DecSmem::W2_STRIDE + (e % (DEC_J / 8)) * 16) =
              __ldg(reinterpret_cast<const uint4*>(p.w2) + e);
        __syncthreads();
        w2_valid = true;
      }
      uint16_t* ybuf = jy + warp * DEC_J;
      for (int li = s + warp * DEC_S; li < nl; li += DEC_S * DEC_WARPS) {
        int4 st = s_state[li];
        const int n = lane0 + li * lstep;
        if (!runnable(st, total)) {
          if (lane == 0) *reinterpret_cast<int4*>(ls_next + (size_t)n * 4) = st;      // carried forward
          continue;
        }
        JointRegs jr = load_counters(p, n, lane);
        float zf[16]; uint4 yv[2];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 v = __ldcg(reinterpret_cast<const float4*>(p.z + (size_t)n * DEC_J + lane * 16) + i);
          zf[4 * i] = v.x; zf[4 * i + 1] = v.y; zf[4 * i + 2] = v.z; zf[4 * i + 3] = v.w;
        }
        load_y1(p, n, st.y, yv, lane);
        bool go = true;
        int budget = p.max_run;
        while (go) {
          joint_y(zf, yv, b1r, ybuf, lane);
          __syncwarp();
          const float logit = __fadd_rn(fc2_chain(w2s, ybuf, lane), s_b2[lane]);
          const int t_before = st.y;
          go = joint_update(p, n, st, jr, logit, lane, total) && --budget > 0;
          ++decisions;
          if (go && st.y != t_before) load_y1(p, n, st.y, yv, lane);
          __syncwarp();
        }
        store_lane(p, ls_next, n, st, jr, lane);
      }
    }
    if (tr) { trow[5] = clock64(); trow[8] = decisions; }
    group_barrier(p, g, epoch, ok, &s_fail);
    par ^= 1;
    if (tr) trow[6] = clock64();
  }

  // ---- epilogue: reference-visible state of lanes that sat at their fixed point while the lock-step loop went on,
  // committed h/c back into the caller's arrays (s_state holds the final states: it was refreshed after the last barrier).
  if (ok && total >= 0) {
    for (int li = s + warp * DEC_S; li < nl; li += DEC_S * DEC_WARPS) {
      const int n = lane0 + li * lstep;
      const int4 st = s_state[li];
      if (lane == 0 && (st.x & LS_FIXED)) p.time_idx[n] += total - st.w;
      if (st.x & LS_CUR) {
        for (int k = lane; k < DEC_P; k += 32) {
          p.hg[0][0][(size_t)n * DEC_P + k] = __ldcg(p.hg[1][0] + (size_t)n * DEC_P + k);
          p.hg[0][1][(size_t)n * DEC_P + k] = __ldcg(p.hg[1][1] + (size_t)n * DEC_P + k);
          p.cg[0][0][(size_t)n * DEC_P + k] = __ldcg(p.cg[1][0] + (size_t)n * DEC_P + k);
          p.cg[0][1][(size_t)n * DEC_P + k] = __ldcg(p.cg[1][1] + (size_t)n * DEC_P + k);
        }
      }
    }
    if (blockIdx.x == 0 && tid == 0 && p.iters_out) *p.iters_out = total;
  }
}

template __global__ void greedy_decode_kernel<true>(const DecParams);
template __global__ void greedy_decode_kernel<false>(const DecParams);

}  // namespace rnnt
