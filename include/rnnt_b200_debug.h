/* rnnt_b200_debug.h — profiling hooks of the DEBUG build of the library only (`make -C rnnt_inference_b200/csrc debug`
 * -> lib/librnnt_b200_dbg.so, compiled with -DRNNT_DEBUG_HOOKS).  The product library librnnt_b200.so exports none of
 * these and its kernels contain none of the trace / switch-off branches.  Used by profiles/trace_*.py. */
#ifndef RNNT_B200_DEBUG_H_
#define RNNT_B200_DEBUG_H_
#ifdef __cplusplus
extern "C" {
#endif
/* Device buffer of T*16 int64: the next recurrent launches record clock64() stamps of CTA (0,0) per timestep (h ready,
 * first operand chunk landed, MMAs issued, accumulator ready, TMEM read, gate math done, stores issued, published ...).
 * NULL switches it off. */
int rnnt_debug_set_recurrent_trace(void* device_buffer);
/* PROFILING ONLY (results become garbage): bit 0 skip the MMAs, 1 skip the gate math, 2 skip the h loads, 3 skip y/c
 * stores, 4 relaxed publish, 5 no proxy fence, 6 no G, 7 no c prefetch, 8 ignore the h_t flags; 0 = normal operation. */
int rnnt_debug_set_recurrent_flags(int flags);
/* PROFILING ONLY (results become garbage): projection GEMM with bit 0 no operand loads, bit 1 no MMAs, bit 2 no result
 * stores, bit 3 every CTA stores all its tiles onto one tile of its own (the result stays in L2: no HBM write-back) — attributes the kernel time
 * to its pipelines (profiles/gemm_attrib.py). */
int rnnt_debug_set_gemm_flags(int flags);
/* Device buffer of max_iters*16 int64 receiving clock64 stamps of CTA 0 per decode iteration. NULL = off. */
int rnnt_debug_set_decode_trace(void* device_buffer, int max_iters);
/* which CTA writes the decode trace (default 0; CTA 20 g + s is CTA s of lane group g) */
int rnnt_debug_set_decode_trace_block(int block);
/* TEST ONLY: CTA 1 of every decode lane group idles `cycles` before each scan of the lane states, i.e. scans while the
 * other CTAs of its group may already be taking the round's decisions.  Results must not change (the lane states are
 * double-buffered by round parity); 0 = off. */
int rnnt_debug_set_decode_skew(int cycles);
#ifdef __cplusplus
}
#endif
#endif
