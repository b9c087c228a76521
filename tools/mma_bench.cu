// Micro-benchmark (not part of the library): cost of one tcgen05.mma kind::f16 (M=128, K=16, SS operands)
// as a function of N and of the operand row width (64-byte SWIZZLE_64B rows vs 128-byte SWIZZLE_128B rows),
// issued back to back by one elected thread on every SM.  Prints cycles per MMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_bench tools/mma_bench.cu && ./mma_bench
#include <cstdio>
#include <cuda_runtime.h>

#include "../medical-image-analysis_b200/csrc/common.cuh"

using namespace b200al;

template <int N, int ROWB>
__global__ void __launch_bounds__(384) mma_loop(int iters, int same_a, long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  // operands: A 128 rows, B N rows (contents irrelevant, zero-filled)
  for (int i = threadIdx.x; i < (4 * 128 + N) * ROWB / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
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
    constexpr uint32_t hi = umma_smem_desc_hi(ROWB);
    const uint32_t a0 = umma_smem_desc_lo(smem_u32(smem));
    const uint32_t b0 = umma_smem_desc_lo(smem_u32(smem + 4 * 128 * ROWB));
    long long t0 = clock64();
    if (elect_one()) {
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          // rotate over 4 A tiles / 4 k-slices so that successive MMAs read different shared memory
          const uint32_t a = a0 + (same_a ? 0 : ((k & 3) * ((128 * ROWB) >> 4))) + ((k >> 2) & 1) * 2;
          umma_bf16(tmem, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, 1u);
        }
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (threadIdx.x == 32) out_cycles[blockIdx.x] = t1 - t0;
  }
  if (warp >= 4) mbar_wait(smem_u32(&bar), 0);  // optional spinning warps (blockDim > 128): mbarrier polling pressure
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

// Tile protocol of the conv kernels: the issuer waits for a free accumulator stage, issues `per_tile`
// MMAs, commits to tfull[stage]; a consumer warp waits tfull and hands the stage back via tempty.
template <int N, int ROWB>
__global__ void __launch_bounds__(128) mma_tiles(int tiles, int per_tile, int nstages, long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t tfull[4], tempty[4];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (4 * 128 + N) * ROWB / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(smem_u32(&tfull[i]), 1); mbar_init(smem_u32(&tempty[i]), 1); }
    fence_barrier_init();
  }
  if (warp == 0) { tmem_alloc(smem_u32(&tmem_slot), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    constexpr uint32_t hi = umma_smem_desc_hi(ROWB);
    const uint32_t a0 = umma_smem_desc_lo(smem_u32(smem));
    const uint32_t b0 = umma_smem_desc_lo(smem_u32(smem + 4 * 128 * ROWB));
    long long t0 = clock64();
    int st = 0; uint32_t ph = 0;
    for (int t = 0; t < tiles; ++t) {
      mbar_wait(smem_u32(&tempty[st]), ph ^ 1);
      tc_fence_after();
      if (elect_one()) {
        for (int k = 0; k < per_tile; ++k) {
          const uint32_t a = a0 + (k & 3) * ((128 * ROWB) >> 4) + ((k >> 2) & 1) * 2;
          umma_bf16(tmem + st * 128, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, k > 0);
        }
        umma_commit(smem_u32(&tfull[st]));
      }
      __syncwarp();
      if (++st == nstages) { st = 0; ph ^= 1; }
    }
    long long t1 = clock64();
    if (threadIdx.x == 32) out_cycles[blockIdx.x] = t1 - t0;
  } else if (warp == 2) {
    int st = 0; uint32_t ph = 0;
    for (int t = 0; t < tiles; ++t) {
      mbar_wait(smem_u32(&tfull[st]), ph);
      tc_fence_after();
      tc_fence_before();
      if (elect_one()) mbar_arrive(smem_u32(&tempty[st]));
      __syncwarp();
      if (++st == nstages) { st = 0; ph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int N, int ROWB>
void run_tiles(int per_tile, int nstages) {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  const int smem = (4 * 128 + N) * ROWB + 1024;
  cudaFuncSetAttribute(mma_tiles<N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int tiles = 1000;
  for (int rep = 0; rep < 2; ++rep) mma_tiles<N, ROWB><<<148, 128, smem>>>(tiles, per_tile, nstages, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < 148; ++i) s += h[i];
  printf("tiles N=%3d per_tile=%2d stages=%d: %.1f cycles/MMA (%s)\n", N, per_tile, nstages,
         s / 148 / ((double)tiles * per_tile), cudaGetErrorString(e));
  cudaFree(d);
}

template <int N, int ROWB>
void run(int same_a, int threads = 128) {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  const int smem = (4 * 128 + N) * ROWB + 1024;
  cudaFuncSetAttribute(mma_loop<N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 2000;
  for (int rep = 0; rep < 2; ++rep) mma_loop<N, ROWB><<<148, threads, smem>>>(iters, same_a, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < 148; ++i) s += h[i];
  printf("N=%3d row=%3dB same_a=%d threads=%d: %.1f cycles/MMA (%s)\n", N, ROWB, same_a, threads, s / 148 / (iters * 8.0),
         cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run_tiles<96, 128>(18, 2); run_tiles<96, 128>(18, 3); run_tiles<96, 128>(18, 4); run_tiles<96, 128>(36, 2);
  run_tiles<96, 128>(6, 2); run_tiles<128, 128>(72, 2); run_tiles<64, 128>(72, 2);
  run<32, 128>(0); run<64, 128>(0); run<96, 128>(0); run<128, 128>(0); run<192, 128>(0); run<256, 128>(0);
  run<32, 64>(0); run<64, 64>(0); run<96, 64>(0); run<128, 64>(0); run<192, 64>(0); run<256, 64>(0);
  run<64, 128>(1); run<96, 128>(1); run<128, 128>(1); run<256, 128>(1);
  run<96, 128>(0, 384); run<128, 128>(0, 384); run<64, 128>(0, 384);
  return 0;
}
