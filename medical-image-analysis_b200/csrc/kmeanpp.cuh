// KMeanSelector pieces (src/activelearning/kmean_selector.py:19-196): per-row standardisation of the embeddings and
// scikit-learn's k-means++ seeding (`sklearn.cluster.kmeans_plusplus`, 1.9.0, `_kmeans_plusplus`) with sample weights.
//
// The random numbers are drawn on the host from numpy's global RandomState in sklearn's order (one `choice` for the
// first centre, `uniform(size=trials)` per further centre) and handed in; everything that touches the n x D matrix
// runs here.  Per centre c >= 1:
//   1. kmpp_sample_kernel   cum = cumsum_f32(w * closest) strictly SEQUENTIAL in fp32 (numpy's accumulate order: the
//                           candidate indices depend on its rounding), candidates = searchsorted(cum, u * pot), clipped.
//                           The chain runs in one thread over shared-memory chunks the rest of the CTA loads and scans.
//   2. kmpp_dist_kernel     d[t][i] = min(closest[i], f32(max(xx_i + cc_t - 2 x_i.c_t, 0) in fp64)) for the <= 16
//                           candidates, one warp per row, candidates resident in shared memory as doubles; weighted
//                           potentials in fp64 per CTA.  (The vector fp64 rate is what bounds this kernel.)
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
  __shared__ __align__(16) float buf[kKmppChunk];
  __shared__ int cnt[kKmppMaxTrials];
  __shared__ float carry_s;
  // searchsorted compares (double)cum[i] < u*pot; for an fp32 cum[i] that is cum[i] < (u*pot rounded UP to fp32),
  // which keeps the fp64 pipe out of the scan
  float r[kKmppMaxTrials];
  const double p = (double)pot[0];
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t) r[t] = t < trials ? __double2float_ru(u[t] * p) : 0.f;
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
      // the dependent fp32 add (4 cycles) is the floor: 128-bit shared-memory accesses (12 instructions per 8
      // elements instead of 24) and the next 8 elements fetched before the current adds (hides the LDS latency)
      float4 a, b;
      if (m >= 8) {
        a = *reinterpret_cast<const float4*>(buf);
        b = *reinterpret_cast<const float4*>(buf + 4);
      }
      for (; j + 8 <= m; j += 8) {
        float4 na = a, nb = b;
        if (j + 16 <= m) {
          na = *reinterpret_cast<const float4*>(buf + j + 8);
          nb = *reinterpret_cast<const float4*>(buf + j + 12);
        }
        c = __fadd_rn(c, a.x); a.x = c;
        c = __fadd_rn(c, a.y); a.y = c;
        c = __fadd_rn(c, a.z); a.z = c;
        c = __fadd_rn(c, a.w); a.w = c;
        c = __fadd_rn(c, b.x); b.x = c;
        c = __fadd_rn(c, b.y); b.y = c;
        c = __fadd_rn(c, b.z); b.z = c;
        c = __fadd_rn(c, b.w); b.w = c;
        *reinterpret_cast<float4*>(buf + j) = a;
        *reinterpret_cast<float4*>(buf + j + 4) = b;
        a = na;
        b = nb;
      }
      for (; j < m; ++j) {
        c = __fadd_rn(c, buf[j]);
        buf[j] = c;
      }
      carry_s = c;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < m; j += blockDim.x) {
      const float v = buf[j];
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

// xx[i] = sum_k x_ik^2 in fp64 (once per call)
template <int NV>
__global__ void __launch_bounds__(256) kmpp_norms_kernel(const float* __restrict__ X, int n, double* __restrict__ xx) {
  constexpr int D = NV * 128;
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp_global; row < n; row += nwarps) {
    float4 x[NV];
    load_row<NV>(X + (size_t)row * D, lane, x);
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      acc = fma((double)x[j].x, (double)x[j].x, acc); acc = fma((double)x[j].y, (double)x[j].y, acc);
      acc = fma((double)x[j].z, (double)x[j].z, acc); acc = fma((double)x[j].w, (double)x[j].w, acc);
    }
    acc = warp_sum_f64(acc);
    if (lane == 0) xx[row] = acc;
  }
}

// grid-stride over rows, one warp per row; dynamic smem = trials * D doubles (candidate rows, converted once per CTA).
// d = xx + cc - 2 x.c evaluated in fp64 - sklearn's own formula (`_euclidean_distances_upcast`) and half the fp64
// work of summing (x - c)^2; a row's distance to itself is forced to 0.
template <int NV>
__global__ void __launch_bounds__(256)
kmpp_dist_kernel(const float* __restrict__ X, const double* __restrict__ xx, const float* __restrict__ w, int n,
                 const int* __restrict__ cand, int trials, const float* __restrict__ closest, int first,
                 float* __restrict__ cand_d, double* __restrict__ part) {
  extern __shared__ double kc[];  // [trials][D]
  constexpr int D = NV * 128;
  __shared__ double red[8][kKmppMaxTrials];
  __shared__ double cc[kKmppMaxTrials];
  __shared__ int cid[kKmppMaxTrials];
  for (int i = threadIdx.x; i < trials * D; i += blockDim.x) kc[i] = (double)X[(size_t)cand[i / D] * D + (i % D)];
  if (threadIdx.x < trials) {
    cid[threadIdx.x] = cand[threadIdx.x];
    cc[threadIdx.x] = xx[cand[threadIdx.x]];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  double pot[kKmppMaxTrials];
#pragma unroll
  for (int t = 0; t < kKmppMaxTrials; ++t) pot[t] = 0.0;
  for (int row = warp_global; row < n; row += nwarps) {
    float4 xf[NV];
    load_row<NV>(X + (size_t)row * D, lane, xf);
    double x[NV][4];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      x[j][0] = (double)xf[j].x; x[j][1] = (double)xf[j].y; x[j][2] = (double)xf[j].z; x[j][3] = (double)xf[j].w;
    }
    const float cur = first ? CUDART_INF_F : closest[row];
    const double wr = (double)w[row], xr = xx[row];
#pragma unroll
    for (int t = 0; t < kKmppMaxTrials; ++t) {
      if (t < trials) {
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;  // four independent chains: the fp64 FMA latency is long
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const double2 c01 = *reinterpret_cast<const double2*>(kc + t * D + (j * 32 + lane) * 4);
          const double2 c23 = *reinterpret_cast<const double2*>(kc + t * D + (j * 32 + lane) * 4 + 2);
          a0 = fma(x[j][0], c01.x, a0); a1 = fma(x[j][1], c01.y, a1);
          a2 = fma(x[j][2], c23.x, a2); a3 = fma(x[j][3], c23.y, a3);
        }
        double acc = warp_sum_f64((a0 + a1) + (a2 + a3));
        double d = xr + cc[t] - 2.0 * acc;
        if (row == cid[t] || d < 0.0) d = 0.0;
        const float m = fminf(cur, (float)d);
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

// <<<1, 32 * trials>>>: warp t reduces the partials of trial t (lane-strided sums, then a fixed shuffle tree)
__global__ void kmpp_pick_kernel(const double* __restrict__ part, int blocks, const int* __restrict__ cand, int trials,
                                 float* __restrict__ pot, int* __restrict__ best, long long* __restrict__ pick_out) {
  __shared__ float pots[kKmppMaxTrials];
  const int t = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double s = 0.0;
  for (int b = lane; b < blocks; b += 32) s += part[(size_t)t * blocks + b];
  s = warp_sum_f64(s);
  if (lane == 0) pots[t] = (float)s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int b = 0;
    for (int q = 1; q < trials; ++q)
      if (pots[q] < pots[b]) b = q;  // np.argmin: first minimum
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
