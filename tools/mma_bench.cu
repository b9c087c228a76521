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
__global__ void __launch_bounds__(128) mma_tiles(int tiles, int per_tile, int nstages, long long* out_cycles, int mode = 0) {
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
    if (mode & 32) {  // one elect for the whole kernel: the lane never leaves the branch; 64: unrolled, constant descriptors
      if (elect_one()) {
        for (int t = 0; t < tiles; ++t) {
          const uint32_t d = tmem + (t & 1) * 128;
          if (mode & 64) {
#pragma unroll
            for (int k = 0; k < 18; ++k) {
              const uint32_t a = a0 + (k & 3) * ((128 * ROWB) >> 4) + ((k >> 2) & 1) * 2;
              umma_bf16(d, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, k > 0);
            }
          } else {
            for (int k = 0; k < per_tile; ++k) {
              const uint32_t a = a0 + (k & 3) * ((128 * ROWB) >> 4) + ((k >> 2) & 1) * 2;
              umma_bf16(d, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, k > 0);
            }
          }
        }
        umma_commit(smem_u32(&tfull[0]));
      }
      tiles = 0;
    }
    if ((mode & 128) && !(mode & 32)) {  // elect per tile, but the 18 MMAs unrolled with constant descriptors
      for (int t = 0; t < tiles; ++t) {
        if (elect_one()) {
          const uint32_t d = tmem + (t & 1) * 128;
#pragma unroll
          for (int k = 0; k < 18; ++k) {
            const uint32_t a = a0 + (k & 3) * ((128 * ROWB) >> 4) + ((k >> 2) & 1) * 2;
            umma_bf16(d, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, k > 0);
          }
          if (t == tiles - 1) umma_commit(smem_u32(&tfull[0]));
        }
        __syncwarp();
      }
      tiles = 0;
    }
    const int tiles_all = tiles;
    for (int t = 0; t < tiles; ++t) {
      // mode bits (where does the per-tile bubble come from?): 1 = always accumulate, 2 = one accumulator address,
      // 4 = no tempty wait, 8 = no fence, 16 = no commit (one at the very end)
      if (!(mode & 4)) mbar_wait(smem_u32(&tempty[st]), ph ^ 1);
      if (!(mode & 8)) tc_fence_after();
      if (elect_one()) {
        const uint32_t d = tmem + ((mode & 2) ? 0 : st * 128);
        for (int k = 0; k < per_tile; ++k) {
          const uint32_t a = a0 + (k & 3) * ((128 * ROWB) >> 4) + ((k >> 2) & 1) * 2;
          umma_bf16(d, make_desc(a, hi), make_desc(b0 + ((k >> 2) & 1) * 2, hi), idesc, (mode & 1) ? 1u : (k > 0));
        }
        if (!(mode & 16) || t == tiles - 1) umma_commit(smem_u32(&tfull[st]));
      }
      __syncwarp();
      if (++st == nstages) { st = 0; ph ^= 1; }
    }
    if (mode & (32 | 128)) mbar_wait(smem_u32(&tfull[0]), 0);  // all MMAs retired
    long long t1 = clock64();
    if (threadIdx.x == 32) out_cycles[blockIdx.x] = t1 - t0;
    (void)tiles_all;
  } else if (warp == 2) {
    int st = 0; uint32_t ph = 0;
    for (int t = 0; t < tiles && !(mode & (4 | 32 | 128)); ++t) {
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

// k-block protocol of the generic conv kernel: a producer warp refills ring slots (here: just arrives),
// the issuer waits full[s], issues `per_blk` MMAs into ONE accumulator, commits to empty[s].
// variant 0: whole warp loops, elect per block (current kernels); 1: one elected thread runs the whole loop;
// 2: like 1, but the full-barrier of the next block is polled before the last two MMAs of this block.
template <int N, int ROWB>
__global__ void __launch_bounds__(128) mma_kblocks(int blocks, int per_blk, int variant, long long* out_cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[4], empty[4], done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (4 * 128 + N) * ROWB / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(smem_u32(&full[i]), 1); mbar_init(smem_u32(&empty[i]), 1); }
    mbar_init(smem_u32(&done), 1);
    fence_barrier_init();
  }
  if (warp == 0) { tmem_alloc(smem_u32(&tmem_slot), 256); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  constexpr uint32_t idesc = umma_idesc_bf16(128, N);
  constexpr uint32_t hi = umma_smem_desc_hi(ROWB);
  const uint32_t a0 = umma_smem_desc_lo(smem_u32(smem));
  const uint32_t b0 = umma_smem_desc_lo(smem_u32(smem + 4 * 128 * ROWB));
  if (warp == 2) {  // producer
    if (threadIdx.x == 64 && (variant < 3 || variant == 9)) {
      int st = 0; uint32_t ph = 0;
      for (int b = 0; b < blocks; ++b) {
        mbar_wait(smem_u32(&empty[st]), ph ^ 1);
        mbar_arrive(smem_u32(&full[st]));
        if (++st == 4) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    long long t0 = clock64();
    if (variant == 9) {
      // software-pipelined bookkeeping: the next block's barrier is polled and its descriptors are built
      // between the two halves of this block's MMAs, so no long gap ever separates two MMAs
      int st = 0; uint32_t ph = 0;
      bool ready = false;
      uint64_t da[4], db[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) { da[k] = make_desc(a0 + k * 2, hi); db[k] = make_desc(b0 + k * 2, hi); }
      for (int b = 0; b < blocks; ++b) {
        if (!ready) mbar_wait(smem_u32(&full[st]), ph);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 6; ++k) umma_bf16(tmem, da[k & 3], db[k & 3], idesc, 1u);
        }
        int nst = st + 1; uint32_t nph = ph;
        if (nst == 4) { nst = 0; nph ^= 1; }
        ready = (b + 1 < blocks) && mbar_try_wait(smem_u32(&full[nst]), nph);
        uint64_t na[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) na[k] = make_desc(a0 + nst * ((128 * ROWB) >> 4) + k * 2, hi);
        if (elect_one()) {
#pragma unroll
          for (int k = 6; k < 12; ++k) umma_bf16(tmem, da[k & 3], db[k & 3], idesc, 1u);
          umma_commit(smem_u32(&empty[st]));
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) da[k] = na[k];
        st = nst; ph = nph;
      }
      if (elect_one()) umma_commit(smem_u32(&done));
      __syncwarp();
    } else if (variant >= 3) {
      // 3: commit per block only; 4: (always-complete) wait + fence per block only; 5: elect per block only;
      // 6: nothing per block (12 unrolled MMAs, one elect at the top)
      int st = 0;
      const bool lead = elect_one();
      for (int b = 0; b < blocks; ++b) {
        if (variant == 4) { mbar_wait(smem_u32(&full[st]), 1); tc_fence_after(); }
        const uint32_t a = a0 + st * ((128 * ROWB) >> 4);
        if (variant >= 7) {
          // descriptors built in uniform control flow, outside the elected branch
          uint64_t da[4], db[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) { da[k] = make_desc(a + k * 2, hi); db[k] = make_desc(b0 + k * 2, hi); }
          if (variant == 8) { mbar_wait(smem_u32(&full[st]), 1); tc_fence_after(); }
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 12; ++k) umma_bf16(tmem, da[k & 3], db[k & 3], idesc, 1u);
            if (variant == 8) umma_commit(smem_u32(&empty[st]));
          }
        } else if (variant == 6 ? lead : elect_one()) {
#pragma unroll
          for (int k = 0; k < 12; ++k)
            umma_bf16(tmem, make_desc(a + (k & 3) * 2, hi), make_desc(b0 + (k & 3) * 2, hi), idesc, 1u);
          if (variant == 3) umma_commit(smem_u32(&empty[st]));
        }
        __syncwarp();
        if (++st == 4) st = 0;
      }
      if (elect_one()) umma_commit(smem_u32(&done));
      __syncwarp();
    } else if (variant == 0) {
      int st = 0; uint32_t ph = 0;
      for (int b = 0; b < blocks; ++b) {
        mbar_wait(smem_u32(&full[st]), ph);
        tc_fence_after();
        const uint32_t a = a0 + st * ((128 * ROWB) >> 4);
        if (elect_one()) {
          for (int k = 0; k < per_blk; ++k)
            umma_bf16(tmem, make_desc(a + (k & 3) * 2, hi), make_desc(b0 + (k & 3) * 2, hi), idesc, 1u);
          umma_commit(smem_u32(&empty[st]));
        }
        __syncwarp();
        if (++st == 4) { st = 0; ph ^= 1; }
      }
      if (elect_one()) umma_commit(smem_u32(&done));
      __syncwarp();
    } else if (elect_one()) {
      int st = 0; uint32_t ph = 0;
      bool ready = false;
      for (int b = 0; b < blocks; ++b) {
        if (!ready) mbar_wait(smem_u32(&full[st]), ph);
        tc_fence_after();
        const uint32_t a = a0 + st * ((128 * ROWB) >> 4);
        int nst = st + 1; uint32_t nph = ph;
        if (nst == 4) { nst = 0; nph ^= 1; }
        ready = false;
        for (int k = 0; k < per_blk; ++k) {
          if (variant == 2 && k == per_blk - 2 && b + 1 < blocks) ready = mbar_try_wait(smem_u32(&full[nst]), nph);
          umma_bf16(tmem, make_desc(a + (k & 3) * 2, hi), make_desc(b0 + (k & 3) * 2, hi), idesc, 1u);
        }
        umma_commit(smem_u32(&empty[st]));
        st = nst; ph = nph;
      }
      umma_commit(smem_u32(&done));
    }
    __syncwarp();
    mbar_wait(smem_u32(&done), 0);
    long long t1 = clock64();
    if (threadIdx.x == 32) out_cycles[blockIdx.x] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int N, int ROWB>
void run_kblocks(int per_blk, int variant) {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  const int smem = (4 * 128 + N) * ROWB + 1024;
  cudaFuncSetAttribute(mma_kblocks<N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int blocks = 2000;
  for (int rep = 0; rep < 2; ++rep) mma_kblocks<N, ROWB><<<148, 128, smem>>>(blocks, per_blk, variant, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < 148; ++i) s += h[i];
  printf("kblocks N=%3d per_blk=%2d variant=%d: %.1f cycles/MMA (%s)\n", N, per_blk, variant,
         s / 148 / ((double)blocks * per_blk), cudaGetErrorString(e));
  cudaFree(d);
}

template <int N, int ROWB>
void run_tiles(int per_tile, int nstages, int mode = 0) {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  const int smem = (4 * 128 + N) * ROWB + 1024;
  cudaFuncSetAttribute(mma_tiles<N, ROWB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int tiles = 1000;
  for (int rep = 0; rep < 2; ++rep) mma_tiles<N, ROWB><<<148, 128, smem>>>(tiles, per_tile, nstages, d, mode);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < 148; ++i) s += h[i];
  printf("tiles N=%3d per_tile=%2d stages=%d mode=%2d: %.1f cycles/MMA (%s)\n", N, per_tile, nstages, mode,
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

int main(int argc, char** argv) {
  if (argc > 1 && argv[1][0] == 'k') {  // k-block protocol, every issue variant
    for (int v = 0; v < 10; ++v) { run_kblocks<128, 128>(12, v); run_kblocks<64, 128>(12, v); run_kblocks<96, 128>(6, v); }
    return 0;
  }
  if (argc > 1) {  // per-tile bubble hunt
    for (int mode : {0, 4, 31, 32, 96, 128}) run_tiles<96, 128>(18, 2, mode);
    return 0;
  }
  for (int v = 7; v < 10; ++v) { run_kblocks<128, 128>(12, v); run_kblocks<64, 128>(12, v); }
  run_kblocks<128, 128>(12, 0); run_kblocks<64, 128>(12, 0);
  run_tiles<96, 128>(18, 2); run_tiles<96, 128>(18, 3); run_tiles<96, 128>(18, 4); run_tiles<96, 128>(36, 2);
  run_tiles<96, 128>(6, 2); run_tiles<128, 128>(72, 2); run_tiles<64, 128>(72, 2);
  run<32, 128>(0); run<64, 128>(0); run<96, 128>(0); run<128, 128>(0); run<192, 128>(0); run<256, 128>(0);
  run<32, 64>(0); run<64, 64>(0); run<96, 64>(0); run<128, 64>(0); run<192, 64>(0); run<256, 64>(0);
  run<64, 128>(1); run<96, 128>(1); run<128, 128>(1); run<256, 128>(1);
  run<96, 128>(0, 384); run<128, 128>(0, 384); run<64, 128>(0, 384);
  return 0;
}
