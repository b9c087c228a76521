// KMeanSelector pieces (src/activelearning/kmean_selector.py:19-196): per-row standardisation of the embeddings and
// scikit-learn's k-means++ seeding (`sklearn.cluster.kmeans_plusplus`, 1.9.0, `_kmeans_plusplus`) with sample weights.
//
// The random numbers are drawn on the host from numpy's global RandomState in sklearn's order (one `choice` for the
// first centre, `uniform(size=trials)` per further centre) and handed in; everything that touches the n x D matrix
// runs here.  Per centre c >= 1:
//   1. kmpp_sample_kernel   cum = cumsum_f32(w * closest) strictly SEQUENTIAL in fp32 (numpy's accumulate order: the
//                           candidate indices depend on its rounding), candidates = searchsorted(cum, u * pot), clipped.
//                           The chain runs in one thread over shared-memory chunks the rest of the CTA loads and scans.
//   2. kmpp_dist_kernel     d[t][i] = min(closest[i], f32(sum_k (x_ik - c_tk)^2 in fp64)) for the <= 16 candidates,
//                           one warp per row, candidates resident in shared memory; weighted potentials in fp64 per CTA.
//   3. kmpp_pick_kernel     potentials reduced in fixed order, rounded to fp32; first minimum wins.
//   4. kmpp_commit_kernel   closest = d[best].
// sklearn evaluates the potentials with an fp32 BLAS dot of CPU-specific summation order; fp64-then-fp32 here agrees
// with it to ~1e-7 relative (see oracle/kmean_ref.py).
#pragma once

#include <math_constants.h>

#include "common.cuh"
#include "kcenter.cuh"

namespace b200al {

constexpr int kKmppMaxTrials = 16;
constexpr int kKmppChunk = 4096;

// (x - mean) / std per row, population std (np.std, ddof = 0); statistics in fp64, rounded to fp32 like numpy's results
__global__ void standardize_rows_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int D) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= n) return;
  const float* r = in + (size_t)row * D;
  double s = 0.0;
  for (int k = lane; k < D; k += 32) s += (double)r[k];
  s = warp_sum_f64(s);
  const float mean = (float)(s / D);
  double v = 0.0;
  for (int k = lane; k < D; k += 32) {
    const double d = (double)__fsub_rn(r[k], mean);
    v = fma(d, d, v);
  }
  v = warp_sum_f64(v);
  const float sd = (float)sqrt(v / D);
  for (int k = lane; k < D; k += 32) out[(size_t)row * D + k] = __fdiv_rn(__fsub_rn(r[k], mean), sd);
}

// <<<1, 1024>>>.  cand[t] = min(#{i : cum[i] < u[t] * pot}, n - 1)
__global__ void __launch_bounds__(1024)
kmpp_sample_kernel(const float* __restrict__ w, const float* __restrict__ closest, int n, const double* __restrict__ u,
                   int trials, const float* __restrict__ pot, int* __restrict__ cand) {
  __shared__ float buf[kKmppChunk];
  __shared__ int cnt[kKmppMaxTrials];
  __shared__ float carry_s;
  double r[kKmppMaxTrials];
  const double p = (double)pot[0];
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t) r[t] = t < trials ? u[t] * p : 0.0;
  int local[kKmppMaxTrials];
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t) local[t] = 0;
  if (threadIdx.x < kKmppMaxTrials) cnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) carry_s = 0.f;
  for (int base = 0; base < n; base += kKmppChunk) {
    const int m = min(kKmppChunk, n - base);
    __syncthreads();
    for (int j = threadIdx.x; j < m; j += blockDim.x) buf[j] = __fmul_rn(w[base + j], closest[base + j]);
    __syncthreads();
    if (threadIdx.x == 0) {
      float c = carry_s;
      int j = 0;
      for (; j + 8 <= m; j += 8) {
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = buf[j + q];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          c = __fadd_rn(c, v[q]);
          buf[j + q] = c;
        }
      }
      for (; j < m; ++j) {
        c = __fadd_rn(c, buf[j]);
        buf[j] = c;
      }
      carry_s = c;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < m; j += blockDim.x) {
      const double v = (double)buf[j];
#pragma unroll
      for (int t = 0; t < kKmppMaxTrials; ++t)
        if (t < trials) local[t] += v < r[t] ? 1 : 0;
    }
  }
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t)
    if (t < trials) {
      int c = local[t];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
      if ((threadIdx.x & 31) == 0 && c) atomicAdd(&cnt[t], c);
    }
  __syncthreads();
  if (threadIdx.x < trials) cand[threadIdx.x] = min(cnt[threadIdx.x], n - 1);
}

// grid-stride over rows, one warp per row; dynamic smem = trials * D floats (candidate rows)
template <int NV>
__global__ void __launch_bounds__(256)
kmpp_dist_kernel(const float* __restrict__ X, const float* __restrict__ w, int n, const int* __restrict__ cand, int trials,
                 const float* __restrict__ closest, int first, float* __restrict__ cand_d, double* __restrict__ part) {
  extern __shared__ float kc[];  // [trials][D]
  constexpr int D = NV * 128;
  __shared__ double red[8][kKmppMaxTrials];
  for (int i = threadIdx.x; i < trials * D; i += blockDim.x) kc[i] = X[(size_t)cand[i / D] * D + (i % D)];
  __syncthreads();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  double pot[kKmppMaxTrials];
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t) pot[t] = 0.0;
  for (int row = warp_global; row < n; row += nwarps) {
    float4 x[NV];
    load_row<NV>(X + (size_t)row * D, lane, x);
    const float cur = first ? CUDART_INF_F : closest[row];
    const double wr = (double)w[row];
#pragma unroll
    for (int t = 0; t < kKmppMaxTrials; ++t) {
      if (t < trials) {
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const float4 c = *reinterpret_cast<const float4*>(kc + t * D + (j * 32 + lane) * 4);
          const double d0 = (double)x[j].x - (double)c.x, d1 = (double)x[j].y - (double)c.y;
          const double d2 = (double)x[j].z - (double)c.z, d3 = (double)x[j].w - (double)c.w;
          acc = fma(d0, d0, acc); acc = fma(d1, d1, acc); acc = fma(d2, d2, acc); acc = fma(d3, d3, acc);
        }
        acc = warp_sum_f64(acc);
        const float m = fminf(cur, (float)acc);
        if (lane == 0) cand_d[(size_t)t * n + row] = m;
        pot[t] = fma(wr, (double)m, pot[t]);
      }
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int t = 0; t < kKmppMaxTrials; ++t) red[wib][t] = pot[t];
  }
  __syncthreads();
  if (threadIdx.x < trials) {
    double s = 0.0;
    for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s += red[q][threadIdx.x];
    part[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = s;
  }
}

// <<<1, 32>>>
__global__ void kmpp_pick_kernel(const double* __restrict__ part, int blocks, const int* __restrict__ cand, int trials,
                                 float* __restrict__ pot, int* __restrict__ best, long long* __restrict__ pick_out) {
  __shared__ float pots[kKmppMaxTrials];
  if ((int)threadIdx.x < trials) {
    double s = 0.0;
    for (int b = 0; b < blocks; ++b) s += part[(size_t)threadIdx.x * blocks + b];
    pots[threadIdx.x] = (float)s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int b = 0;
    for (int t = 1; t < trials; ++t)
      if (pots[t] < pots[b]) b = t;  // np.argmin: first minimum
    pot[0] = pots[b];
    best[0] = b;
    pick_out[0] = cand[b];
  }
}

__global__ void kmpp_commit_kernel(const float* __restrict__ cand_d, const int* __restrict__ best, int n,
                                   float* __restrict__ closest) {
  const float* src = cand_d + (size_t)best[0] * n;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) closest[i] = src[i];
}

__global__ void kmpp_set_first_kernel(int* cand, int first) { cand[0] = first; }

}  // namespace b200al
