// Coreset selection on TWO feature matrices, coreset_fusion="add" (coreset_selector.py:110-157): model embeddings
// plus externally loaded (foundation-model) features, each distance matrix divided by its own GLOBAL sum and mixed
//     final(i, j) = w * D_loaded(i, j) / sum(D_loaded) + (1 - w) * D_feat(i, j) / sum(D_feat)
// before k-center greedy.  The global sums are not argmax-invariant once two matrices are added, so they are
// computed for real (pair_distance_sum_kernel: one pass over all N^2 pairs, fp64 accumulation, never materialised);
// the greedy loop then keeps a running min (fp32) or running sum (fp64) of the fused distance to the growing centre
// set, like csrc/kcenter.cuh.  Element arithmetic follows the reference's fp32 order:
//     fused = fl32( fl32(w_l * fl32(d_l / S_l)) + fl32(w_f * fl32(d_f / S_f)) )
// Feature widths are arbitrary here (loaded features are 200..1024 wide), so rows are read with plain strided loops;
// this path is meant for pools of up to a few 10^4 rows (the reference itself needs three N x N fp32 matrices).
#pragma once

#include <math_constants.h>

#include "common.cuh"
#include "kcenter.cuh"

namespace b200al {

struct FusedFeats {
  const float* e1;  // model embeddings  [n, d1] (normalised rows for cosine)
  const float* e2;  // loaded features   [n, d2]
  int d1, d2, metric;
  float s1, s2;     // global pair-distance sums of each matrix (fp32, like `dist.sum()`)
  float w1, w2;     // 1 - loaded_feature_weight, loaded_feature_weight
};

// sklearn fp32 distance of two rows (same arithmetic as row_distance<> in kcenter.cuh), any width
__device__ __forceinline__ float row_distance_any(const float* __restrict__ x, const float* __restrict__ c, int D,
                                                  int lane, int metric) {
  double acc = 0.0;
  if (metric == METRIC_L2) {
    for (int k = lane; k < D; k += 32) {
      const double d = (double)x[k] - (double)c[k];
      acc = fma(d, d, acc);
    }
    acc = warp_sum_f64(acc);
    return sqrtf((float)acc);
  }
  for (int k = lane; k < D; k += 32) acc = fma((double)x[k], (double)c[k], acc);
  acc = warp_sum_f64(acc);
  return fminf(fmaxf(1.0f - (float)acc, 0.0f), 2.0f);
}

__device__ __forceinline__ float fused_distance(const FusedFeats& f, int row, int centre, int lane) {
  if (row == centre) return 0.f;  // pairwise_distances(X) zeroes its diagonal
  const float dl = row_distance_any(f.e2 + (size_t)row * f.d2, f.e2 + (size_t)centre * f.d2, f.d2, lane, f.metric);
  const float df = row_distance_any(f.e1 + (size_t)row * f.d1, f.e1 + (size_t)centre * f.d1, f.d1, lane, f.metric);
  return __fadd_rn(__fmul_rn(f.w2, __fdiv_rn(dl, f.s2)), __fmul_rn(f.w1, __fdiv_rn(df, f.s1)));
}

// partial[warp_global] = sum over this warp's rows i of sum_j d(i, j); summed in index order by the caller
__global__ void __launch_bounds__(256)
pair_distance_sum_kernel(const float* __restrict__ e, int n, int D, int metric, double* __restrict__ partial) {
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  double s = 0.0;
  for (int i = warp_global; i < n; i += nwarps)
    for (int j = 0; j < n; ++j)
      if (j != i) s += (double)row_distance_any(e + (size_t)i * D, e + (size_t)j * D, D, lane, metric);
  if (lane == 0) partial[warp_global] = s;
}

__global__ void sum_partials_kernel(const double* __restrict__ partial, int count, float* __restrict__ out) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < count; ++i) s += partial[i];
    out[0] = (float)s;
  }
}

// state[i] = min / sum over the L initial centres (rows 0..L-1) of the fused distance
__global__ void __launch_bounds__(256)
kcenter_fused_init_kernel(FusedFeats f, const float* __restrict__ sums, int n, int L, int criteria,
                          void* __restrict__ state) {
  f.s1 = sums[0];
  f.s2 = sums[1];
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp_global; row < n; row += nwarps) {
    float mn = CUDART_INF_F;
    double sm = 0.0;
    for (int j = 0; j < L; ++j) {
      const float d = fused_distance(f, row, j, lane);
      mn = fminf(mn, d);
      sm += (double)d;
    }
    if (lane == 0) {
      if (criteria == CRITERIA_MEAN) ((double*)state)[row] = sm;
      else ((float*)state)[row] = mn;
    }
  }
}

// one greedy step, part 1: fold the newest centre (picks[step-1]) into the state, per-block masked argmax
__global__ void __launch_bounds__(256)
kcenter_fused_step_kernel(FusedFeats f, const float* __restrict__ sums, int n, const long long* __restrict__ picks,
                          int step, int criteria, void* __restrict__ state, double inv_count,
                          const uint8_t* __restrict__ taken, float* __restrict__ part_val, int* __restrict__ part_idx) {
  f.s1 = sums[0];
  f.s2 = sums[1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int rows_per_block = (n + gridDim.x - 1) / gridDim.x;
  const int r_begin = blockIdx.x * rows_per_block;
  const int r_end = min(n, r_begin + rows_per_block);
  const int centre = step > 0 ? (int)picks[step - 1] : -1;
  float bv = -CUDART_INF_F;
  int bi = 0x7fffffff;
  for (int row = r_begin + warp; row < r_end; row += nw) {
    float v;
    if (criteria == CRITERIA_MEAN) {
      double m = ((double*)state)[row];
      if (centre >= 0) {
        m += (double)fused_distance(f, row, centre, lane);
        if (lane == 0) ((double*)state)[row] = m;
      }
      v = (float)(m * inv_count);
    } else {
      float m = ((float*)state)[row];
      if (centre >= 0) {
        m = fminf(m, fused_distance(f, row, centre, lane));
        if (lane == 0) ((float*)state)[row] = m;
      }
      v = m;
    }
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

}  // namespace b200al
