// mma_rate.cu — microbenchmark (profiling aid, not product code): how long does a burst of tcgen05.mma.kind::i8
// instructions take on one SM as a function of N and of the accumulator dependency pattern?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../rnnt_inference_b200/csrc -o mma_rate.bin mma_rate.cu
// Operands are whatever is in shared memory (zeros); only timing matters.  One CTA per SM on `ctas` SMs.
#include <cstdio>
#include "common.cuh"
using namespace rnnt;

template <int N, int PARTS>
__global__ void __launch_bounds__(128, 1) mma_burst(long long* out, int reps, int mmas) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = base, smem_b = base + 8 * 16384, bar = base + 8 * 16384 + 2 * N * 128, slot = bar + 16;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  for (int i = threadIdx.x; i < (8 * 16384 + 2 * N * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(gen)[i] = 0x01010101u;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(slot);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *(volatile uint32_t*)(gen + (slot - base));
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_i8(128, N);
    uint32_t phase = 0;
    long long best = 1ll << 60;
    for (int r = 0; r < reps; ++r) {
      const long long t0 = clock64();
      for (int i = 0; i < mmas; ++i) {
        const int kc = (i >> 2) & 7, k = i & 3;
        const uint64_t ad = umma_desc_sw128(smem_a + kc * 16384) + (uint64_t)(k * 2);
        const uint64_t bd = umma_desc_sw128(smem_b + (kc & 1) * N * 128) + (uint64_t)(k * 2);
        umma_i8(tmem + (uint32_t)(i % PARTS) * N, ad, bd, idesc, i >= PARTS);
      }
      umma_commit(bar);
      while (!mbar_try_wait(bar, phase)) {}
      phase ^= 1;
      const long long dt = clock64() - t0;
      if (dt < best) best = dt;
    }
    out[blockIdx.x] = best;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

template <int N, int PARTS>
void run(const char* name, int ctas, int mmas) {
  long long* d; cudaMalloc(&d, ctas * sizeof(long long));
  const int smem = 8 * 16384 + 2 * N * 128 + 1024 + 64;
  cudaFuncSetAttribute(mma_burst<N, PARTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  mma_burst<N, PARTS><<<ctas, 128, smem>>>(d, 20, mmas);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[160] = {0};
  cudaMemcpy(h, d, ctas * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("%-34s ctas=%3d mmas=%3d : %6lld cycles = %6.1f cycles/MMA, %7.0f MAC/clk/SM  (%s)\n", name, ctas, mmas, mx,
         (double)mx / mmas, 128.0 * N * 32 * mmas / mx, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int ctas : {1, 128}) {
    run<64, 1>("M128 N64  one accumulator", ctas, 32);
    run<64, 4>("M128 N64  4 partial accumulators", ctas, 32);
    run<64, 1>("M128 N64  one accumulator", ctas, 256);
    run<64, 4>("M128 N64  4 partial accumulators", ctas, 256);
    run<128, 1>("M128 N128 one accumulator", ctas, 32);
    run<128, 2>("M128 N128 2 partial accumulators", ctas, 32);
    run<128, 1>("M128 N128 one accumulator", ctas, 256);
    run<256, 1>("M128 N256 one accumulator", ctas, 32);
    run<256, 1>("M128 N256 one accumulator", ctas, 256);
  }
  return 0;
}
