// Micro-benchmark (not part of the library): tcgen05.mma kind::f16, M=128, K=16, with DIFFERENT row widths for
// the two shared-memory operands (A in 128-byte SWIZZLE_128B rows, B in 64- or 128-byte rows) and N up to 256.
// Question answered: does a_hi * [w_hi ; w_lo] as one N=192 MMA with the weights in 64-byte rows beat two
// N=96 MMAs with packed 128-byte rows in the Cin=32 folded kernel?  Prints cycles per MMA on SM 0.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_mix_bench tools/mma_mix_bench.cu && ./mma_mix_bench
#include <cstdio>
#include <cuda_runtime.h>

#include "../medical-image-analysis_b200/csrc/common.cuh"

using namespace b200al;

template <int N, int ROWA, int ROWB>
__global__ void __launch_bounds__(128) mma_loop(int iters, long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  constexpr int kABytes = 4 * 128 * ROWA;
  for (int i = threadIdx.x; i < (kABytes + 256 * ROWB) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(smem_u32(&tmem_slot), 256);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    constexpr uint32_t hia = umma_smem_desc_hi(ROWA);
    constexpr uint32_t hib = umma_smem_desc_hi(ROWB);
    const uint32_t a0 = umma_smem_desc_lo(smem_u32(smem));
    const uint32_t b0 = umma_smem_desc_lo(smem_u32(smem + kABytes));
    long long t0 = clock64();
    if (elect_one()) {
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a = a0 + (k & 3) * ((128 * ROWA) >> 4) + ((k >> 2) & 1) * 2;
          umma_bf16(tmem, make_desc(a, hia), make_desc(b0 + ((k >> 2) & 1) * 2, hib), idesc, 1u);
        }
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (threadIdx.x == 32) out_cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int N, int ROWA, int ROWB>
void run(long long* d) {
  const int iters = 2000, smem = 4 * 128 * ROWA + 256 * ROWB + 1024;
  cudaFuncSetAttribute(mma_loop<N, ROWA, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) mma_loop<N, ROWA, ROWB><<<148, 128, smem>>>(iters, d);
  long long h = 0;
  cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaError_t e = cudaDeviceSynchronize();
  printf("N=%3d A rows %3d B  B rows %3d B: %6.1f cycles per MMA (math %d)%s\n", N, ROWA, ROWB,
         (double)h / (iters * 8), N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  run<96, 128, 128>(d);
  run<96, 128, 64>(d);
  run<96, 64, 64>(d);
  run<192, 128, 128>(d);
  run<192, 128, 64>(d);
  run<192, 64, 64>(d);
  run<64, 128, 128>(d);
  run<128, 128, 128>(d);
  run<128, 128, 64>(d);
  run<256, 128, 128>(d);
  return 0;
}
