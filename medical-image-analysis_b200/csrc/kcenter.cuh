// k-center-greedy coreset selection on embeddings [n, D] fp32 (kernel 4).
// Restates kcenter_greedy (coreset_selector.py:19-51, criteria "min") as a
// running minimum: mind[i] = min_j d(i, centre_j); every step takes the FIRST
// maximum of mind over untaken rows (numpy argmax, :43), marks it, and folds
// its distances into mind.  The N x N matrix (coreset_selector.py:132,152) is
// never built.
//
// Distances reproduce scikit-learn's fp32 results (the reference's
// pairwise_distances, sklearn 1.9.0):
//   l2:     d = sqrt_f32( f32( sum_k (x_k - c_k)^2 accumulated in fp64 ) )
//   cosine: rows pre-normalised in fp32; d = clip(1 - f32(dot in fp64), 0, 2)
// One warp per row, float4 loads, fp64 accumulate, fixed-order shuffles.
// The step is HBM-bound: n*(4*D + 9) bytes per step.
#pragma once

#include <math_constants.h>

#include "common.cuh"

namespace b200al {

enum { METRIC_L2 = 0, METRIC_COSINE = 1 };
enum { CRITERIA_MIN = 0, CRITERIA_MEAN = 1 };

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// NV = D / 128 float4 per lane
template <int NV>
__device__ __forceinline__ void load_row(const float* __restrict__ row, int lane, float4 (&r)[NV]) {
#pragma unroll
  for (int j = 0; j < NV; ++j) r[j] = __ldg(reinterpret_cast<const float4*>(row) + j * 32 + lane);
}

template <int NV>
__device__ __forceinline__ float row_distance(const float4 (&x)[NV], const float4 (&c)[NV], int metric) {
  double acc = 0.0;
  if (metric == METRIC_L2) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const double d0 = (double)x[j].x - (double)c[j].x, d1 = (double)x[j].y - (double)c[j].y;
      const double d2 = (double)x[j].z - (double)c[j].z, d3 = (double)x[j].w - (double)c[j].w;
      acc = fma(d0, d0, acc); acc = fma(d1, d1, acc); acc = fma(d2, d2, acc); acc = fma(d3, d3, acc);
    }
    acc = warp_sum_f64(acc);
    return sqrtf((float)acc);
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    acc = fma((double)x[j].x, (double)c[j].x, acc); acc = fma((double)x[j].y, (double)c[j].y, acc);
    acc = fma((double)x[j].z, (double)c[j].z, acc); acc = fma((double)x[j].w, (double)c[j].w, acc);
  }
  acc = warp_sum_f64(acc);
  return fminf(fmaxf(1.0f - (float)acc, 0.0f), 2.0f);
}

// sklearn normalize(): x / sqrt(sum x^2) in fp32, zero rows untouched
__global__ void normalize_rows_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int D) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const float* r = in + (size_t)warp * D;
  float s = 0.f;
  for (int k = lane; k < D; k += 32) s = fmaf(r[k], r[k], s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  float nrm = sqrtf(s);
  if (nrm == 0.f) nrm = 1.f;
  for (int k = lane; k < D; k += 32) out[(size_t)warp * D + k] = r[k] / nrm;
}

// mind[i] = min(mind_in[i] or +inf, min_j d(row i, centre j)), j < L.  Each warp keeps
// ROWS rows in registers and streams the L centres (L2-resident) past them.
template <int NV, int ROWS>
__global__ void __launch_bounds__(256)
kcenter_init_kernel(const float* __restrict__ emb, int n, const float* __restrict__ centres, int L, int metric,
                    float* __restrict__ mind, int accumulate) {
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  constexpr int D = NV * 128;
  for (int r0 = warp_global * ROWS; r0 < n; r0 += nwarps * ROWS) {
    float4 x[ROWS][NV];
    float best[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int row = min(r0 + r, n - 1);
      load_row<NV>(emb + (size_t)row * D, lane, x[r]);
      best[r] = accumulate ? mind[row] : CUDART_INF_F;
    }
    for (int j = 0; j < L; ++j) {
      float4 c[NV];
      load_row<NV>(centres + (size_t)j * D, lane, c);
#pragma unroll
      for (int r = 0; r < ROWS; ++r) best[r] = fminf(best[r], row_distance<NV>(x[r], c, metric));
    }
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < ROWS; ++r)
        if (r0 + r < n) mind[r0 + r] = best[r];
    }
  }
}

// One greedy step, part 1: fold the newest centre into mind (if do_update), then
// per-block masked argmax (largest mind, lowest row on ties).
template <int NV>
__global__ void __launch_bounds__(256)
kcenter_step_kernel(const float* __restrict__ emb, int n, const float* __restrict__ centre, int do_update, int metric,
                    float* __restrict__ mind, const uint8_t* __restrict__ taken, float* __restrict__ part_val,
                    int* __restrict__ part_idx) {
  constexpr int D = NV * 128;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int rows_per_block = (n + gridDim.x - 1) / gridDim.x;
  const int r_begin = blockIdx.x * rows_per_block;
  const int r_end = min(n, r_begin + rows_per_block);
  float4 c[NV];
  if (do_update) load_row<NV>(centre, lane, c);
  float bv = -CUDART_INF_F;
  int bi = 0x7fffffff;
  for (int row = r_begin + warp; row < r_end; row += nw) {
    float m = mind[row];
    if (do_update) {
      float4 x[NV];
      load_row<NV>(emb + (size_t)row * D, lane, x);
      m = fminf(m, row_distance<NV>(x, c, metric));
      if (lane == 0) mind[row] = m;
    }
    if (!taken[row] && (m > bv)) { bv = m; bi = row; }  // rows ascend: strict > keeps the first max
  }
  __shared__ float sv[32];
  __shared__ int si[32];
  if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < nw; ++w)
      if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
    part_val[blockIdx.x] = bv;
    part_idx[blockIdx.x] = bi;
  }
}

// ---- coreset_criteria == "mean" (coreset_selector.py:35-36): the score of an unlabeled row is the MEAN
// distance to all current centres.  Kept as a running fp64 sum per row (the reference re-averages an fp32
// matrix slice every step); the divisor is the same for every row, so it only scales the reported value.
template <int NV, int ROWS>
__global__ void __launch_bounds__(256)
kcenter_init_mean_kernel(const float* __restrict__ emb, int n, const float* __restrict__ centres, int L, int metric,
                         double* __restrict__ msum, int accumulate) {
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  constexpr int D = NV * 128;
  for (int r0 = warp_global * ROWS; r0 < n; r0 += nwarps * ROWS) {
    float4 x[ROWS][NV];
    double sum[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int row = min(r0 + r, n - 1);
      load_row<NV>(emb + (size_t)row * D, lane, x[r]);
      sum[r] = accumulate ? msum[row] : 0.0;
    }
    for (int j = 0; j < L; ++j) {
      float4 c[NV];
      load_row<NV>(centres + (size_t)j * D, lane, c);
#pragma unroll
      for (int r = 0; r < ROWS; ++r) sum[r] += (double)row_distance<NV>(x[r], c, metric);
    }
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < ROWS; ++r)
        if (r0 + r < n) msum[r0 + r] = sum[r];
    }
  }
}

template <int NV>
__global__ void __launch_bounds__(256)
kcenter_step_mean_kernel(const float* __restrict__ emb, int n, const float* __restrict__ centre, int do_update,
                         int metric, double* __restrict__ msum, double inv_count, const uint8_t* __restrict__ taken,
                         float* __restrict__ part_val, int* __restrict__ part_idx) {
  constexpr int D = NV * 128;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int rows_per_block = (n + gridDim.x - 1) / gridDim.x;
  const int r_begin = blockIdx.x * rows_per_block;
  const int r_end = min(n, r_begin + rows_per_block);
  float4 c[NV];
  if (do_update) load_row<NV>(centre, lane, c);
  float bv = -CUDART_INF_F;
  int bi = 0x7fffffff;
  for (int row = r_begin + warp; row < r_end; row += nw) {
    double m = msum[row];
    if (do_update) {
      float4 x[NV];
      load_row<NV>(emb + (size_t)row * D, lane, x);
      m += (double)row_distance<NV>(x, c, metric);
      if (lane == 0) msum[row] = m;
    }
    const float v = (float)(m * inv_count);
    if (!taken[row] && (v > bv)) { bv = v; bi = row; }
  }
  __shared__ float sv[32];
  __shared__ int si[32];
  if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < nw; ++w)
      if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
    part_val[blockIdx.x] = bv;
    part_idx[blockIdx.x] = bi;
  }
}

// part 2 (one block): global argmax over block partials, record the pick, mark
// it taken, publish its row as the next centre.
__global__ void kcenter_pick_kernel(const float* __restrict__ part_val, const int* __restrict__ part_idx, int nparts,
                                    const float* __restrict__ emb, int D, uint8_t* __restrict__ taken,
                                    float* __restrict__ centre, long long* __restrict__ picks,
                                    float* __restrict__ pick_vals, int step) {
  __shared__ float sv[256];
  __shared__ int si[256];
  float bv = -CUDART_INF_F;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
    const float v = part_val[i];
    const int ix = part_idx[i];
    if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
  }
  sv[threadIdx.x] = bv;
  si[threadIdx.x] = bi;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const float v = sv[threadIdx.x + o];
      const int ix = si[threadIdx.x + o];
      if (v > sv[threadIdx.x] || (v == sv[threadIdx.x] && ix < si[threadIdx.x])) {
        sv[threadIdx.x] = v;
        si[threadIdx.x] = ix;
      }
    }
    __syncthreads();
  }
  const int best = si[0];
  if (best == 0x7fffffff) {  // nothing left to pick
    if (threadIdx.x == 0) { picks[step] = -1; if (pick_vals) pick_vals[step] = -CUDART_INF_F; }
    return;
  }
  if (threadIdx.x == 0) {
    picks[step] = best;
    if (pick_vals) pick_vals[step] = sv[0];
    taken[best] = 1;
  }
  for (int k = threadIdx.x; k < D; k += blockDim.x) centre[k] = emb[(size_t)best * D + k];
}

}  // namespace b200al
