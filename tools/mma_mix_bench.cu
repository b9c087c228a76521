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

// The folded Cin = 64 tile as issued: 12 stacked MMAs (N = 192, D columns 0..191) then 12 of N = 96 into the first half,
// accumulate = 0 on the first, alternating between two accumulators; no barriers, no commits inside the loop.
// seq: 0 = as above, 1 = all 24 with N = 192, 2 = all 24 with N = 96, 3 = N alternating every MMA
__global__ void __launch_bounds__(128) mma_fold_tile(int tiles, int seq, long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  constexpr int kABytes = 4 * 128 * 128;
  for (int i = threadIdx.x; i < (kABytes + 6 * 96 * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&tmem_slot), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 1) {
    constexpr uint32_t i192 = umma_idesc_bf16(128, 192), i96 = umma_idesc_bf16(128, 96);
    constexpr uint32_t hi = umma_smem_desc_hi(128);
    const uint32_t a0 = umma_smem_desc_lo(smem_u32(smem));
    const uint32_t b0 = umma_smem_desc_lo(smem_u32(smem + kABytes));
    long long t0 = clock64();
    if (elect_one()) {
      for (int t = 0; t < tiles; ++t) {
        const uint32_t d = tmem + (t & 1) * 256;
        if (seq >= 4) {
          // the kernel's exact operand walk: A = halo slot (6 x 32 px rows of 128 B), filter row r = +32 rows,
          // k-slice = +32 B; B = resident weights [r][hi 96 rows | lo 96 rows]; hi slot then lo slot.
          // seq 5: the same with the k loop outside the r loop (consecutive MMAs touch different rows)
#pragma unroll
          for (int ph = 0; ph < 2; ++ph)
#pragma unroll
            for (int m = 0; m < 12; ++m) {
              const int r = seq == 4 ? m / 4 : m % 3, k = seq == 4 ? m % 4 : m / 3;
              const uint32_t a = a0 + ((ph * 6 * 32 * 128 + r * 32 * 128 + k * 32) >> 4);
              const uint32_t b = b0 + ((r * 2 * 96 * 128 + k * 32) >> 4);
              umma_bf16(d, make_desc(a, hi), make_desc(b, hi), ph == 0 ? i192 : i96, (ph | m) > 0);
            }
          continue;
        }
#pragma unroll
        for (int m = 0; m < 24; ++m) {
          const uint32_t a = a0 + (m & 3) * ((128 * 128) >> 4) + ((m >> 2) & 3) * 2;
          const bool wide = seq == 0 ? m < 12 : (seq == 1 ? true : (seq == 2 ? false : (m & 1) == 0));
          umma_bf16(d, make_desc(a, hi), make_desc(b0 + ((m >> 2) & 3) * 2, hi), wide ? i192 : i96, m > 0);
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
  if (warp == 0) tmem_dealloc(tmem, 512);
}

void run_fold_tile(long long* d, int seq) {
  const int tiles = 500, smem = 4 * 128 * 128 + 6 * 96 * 128 + 1024;
  cudaFuncSetAttribute(mma_fold_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) mma_fold_tile<<<148, 128, smem>>>(tiles, seq, d);
  long long h = 0;
  cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaError_t e = cudaDeviceSynchronize();
  printf("fold tile seq=%d: %7.1f cycles per tile of 24 MMAs (12 x 96 + 12 x 56 = 1824 expected for seq 0) %s\n", seq,
         (double)h / tiles, e == cudaSuccess ? "" : cudaGetErrorString(e));
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
  for (int seq = 0; seq < 6; ++seq) run_fold_tile(d, seq);
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
