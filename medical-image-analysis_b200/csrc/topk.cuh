// Rank & cut (entropy_selector.py:76-82): the k best of n scores, descending,
// lower position first on ties (the declared tie-break; torch.sort there is
// unstable).  Exact: 64-bit keys = (order-preserving score bits, position) are
// unique, an 8-pass MSB radix SELECT finds the k-th key, survivors are
// compacted and ordered by counting rank (k^2 compares, k <= a few thousand).
// Integer atomics only -> deterministic output.
#pragma once

#include "common.cuh"

namespace b200al {

struct TopkState {
  unsigned long long prefix;  // bytes of the k-th key decided so far (MSB first)
  unsigned int remaining;     // rank still to resolve inside the current prefix bucket
  unsigned int count;         // compaction cursor
  unsigned int hist[256];
};

// smaller key == better (higher score, then lower position)
__device__ __forceinline__ unsigned long long topk_key(float score, unsigned int idx) {
  unsigned int u = __float_as_uint(score);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // ascending-order bits
  return ((unsigned long long)(~u) << 32) | idx;
}
__device__ __forceinline__ float topk_key_score(unsigned long long key) {
  unsigned int u = ~(unsigned int)(key >> 32);
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}

__global__ void topk_keys_kernel(const float* __restrict__ scores, int n, unsigned long long* __restrict__ keys,
                                 TopkState* st, unsigned int k) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    keys[i] = topk_key(scores[i], (unsigned int)i);
  if (blockIdx.x == 0) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) st->hist[i] = 0;
    if (threadIdx.x == 0) { st->prefix = 0ull; st->remaining = k; st->count = 0; }
  }
}

// histogram of byte `pass` (0 = most significant) among keys matching the prefix so far
__global__ void topk_hist_kernel(const unsigned long long* __restrict__ keys, int n, TopkState* st, int pass) {
  __shared__ unsigned int h[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const int shift = 56 - 8 * pass;
  const unsigned long long prefix = st->prefix;
  const unsigned long long himask = pass == 0 ? 0ull : (~0ull << (shift + 8));
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned long long key = keys[i];
    if ((key & himask) == prefix) atomicAdd(&h[(key >> shift) & 0xff], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x)
    if (h[i]) atomicAdd(&st->hist[i], h[i]);
}

// one thread: pick the bucket holding the `remaining`-th smallest key
__global__ void topk_scan_kernel(TopkState* st, int pass) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int shift = 56 - 8 * pass;
  unsigned int rem = st->remaining, cum = 0;
  int b = 0;
  for (; b < 256; ++b) {
    const unsigned int c = st->hist[b];
    if (cum + c >= rem) break;
    cum += c;
  }
  if (b == 256) b = 255;
  st->prefix |= ((unsigned long long)b << shift);
  st->remaining = rem - cum;
  for (int i = 0; i < 256; ++i) st->hist[i] = 0;
}

__global__ void topk_compact_kernel(const unsigned long long* __restrict__ keys, int n, TopkState* st,
                                    unsigned long long* __restrict__ cand, unsigned int k) {
  const unsigned long long kth = st->prefix;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const unsigned long long key = keys[i];
    if (key <= kth) {
      const unsigned int slot = atomicAdd(&st->count, 1u);
      if (slot < k) cand[slot] = key;
    }
  }
}

// rank by counting (keys unique) -> sorted output
__global__ void topk_rank_kernel(const unsigned long long* __restrict__ cand, unsigned int k, long long base_pos,
                                 float* __restrict__ out_scores, long long* __restrict__ out_pos) {
  __shared__ unsigned long long tile[256];
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long mine = i < k ? cand[i] : ~0ull;
  unsigned int rank = 0;
  for (unsigned int t0 = 0; t0 < k; t0 += 256) {
    tile[threadIdx.x] = (t0 + threadIdx.x < k) ? cand[t0 + threadIdx.x] : ~0ull;
    __syncthreads();
    const unsigned int lim = min(256u, k - t0);
    for (unsigned int j = 0; j < lim; ++j) rank += (tile[j] < mine);
    __syncthreads();
  }
  if (i < k) {
    out_scores[rank] = topk_key_score(mine);
    out_pos[rank] = base_pos + (long long)(mine & 0xffffffffull);
  }
}

}  // namespace b200al
