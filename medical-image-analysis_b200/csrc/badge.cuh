// BADGE gradient embedding (src/activelearning/badge_selector.py:19-104): per pool image, the gradient of
// CE(logits, pseudo-labels) + Dice(logits, pseudo-labels) w.r.t. the 1x1 head weight `decoder.seg_output.weight`
// [K, 32], with pseudo-labels = argmax softmax (badge_selector.py:79-80, 91-95) and the losses as al_train builds
// them (al_trainer.py:782-798; dice_loss.py:32-76).  The reference runs a full autograd backward per image; only the
// head's weight gradient is kept, which has a closed form over the last 32-channel feature map:
//     p = softmax(z), y = onehot(argmax p), P = H*W
//     I_c = sum p_c y_c, U_c = sum p_c + sum y_c + s            (per image, dice classes c; C of them)
//     q_c = (-2 y_c / U_c + (2 I_c + s) / U_c^2) / C            (d dice / d p_c; 0 for a skipped background class)
//     g_k = (p_k - y_k) / P + p_k (q_k - sum_j p_j q_j)         (d loss / d z_k)
//     dL/dW[k][c] = sum_px g_k(px) feat_c(px)
// Two passes over the feature map (bf16 hi/lo, 128 B per pixel): badge_stats_kernel accumulates (I, sum p, sum y),
// badge_grad_kernel the K x 32 outer-product sums; logits are recomputed from the features on the fly.  All
// reductions are fixed-order (thread -> warp -> block -> per-image partials summed in index order in fp64).
#pragma once

#include "common.cuh"
#include "uncertainty.cuh"

namespace b200al {

constexpr int kBadgeBlocks = 64;   // CTAs per image
constexpr int kBadgeMaxK = 4;      // K*32 accumulators per thread

template <int K>
__device__ __forceinline__ void badge_pixel(const __nv_bfloat16* __restrict__ src, int hi_off, int lo_off, int use_lo,
                                            const float* sw, const float* sb, float (&f)[32], float (&p)[K], int& y) {
  const uint4* ph = reinterpret_cast<const uint4*>(src + hi_off);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint4 u = __ldg(ph + j);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      f[j * 8 + e * 2] = bf16lo_to_float(w[e]);
      f[j * 8 + e * 2 + 1] = bf16hi_to_float(w[e]);
    }
  }
  if (use_lo) {
    const uint4* pl = reinterpret_cast<const uint4*>(src + lo_off);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint4 u = __ldg(pl + j);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        f[j * 8 + e * 2] += bf16lo_to_float(w[e]);
        f[j * 8 + e * 2 + 1] += bf16hi_to_float(w[e]);
      }
    }
  }
  float z[K], m = -3.4e38f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    float a = sb[k];
#pragma unroll
    for (int c = 0; c < 32; ++c) a = fmaf(f[c], sw[k * 32 + c], a);
    z[k] = a;
    m = fmaxf(m, a);
  }
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    p[k] = expf(z[k] - m);
    sum += p[k];
  }
  const float inv = 1.0f / sum;
  y = 0;
  float best = -1.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    p[k] *= inv;
    if (p[k] > best) {  // torch.argmax: first maximum
      best = p[k];
      y = k;
    }
  }
}

// grid (kBadgeBlocks, n); stats_partial [n][kBadgeBlocks][3K] = (I_c, sum p_c, sum y_c)
template <int K>
__global__ void __launch_bounds__(256)
badge_stats_kernel(const __nv_bfloat16* __restrict__ feat, int pitch, int hi_off, int lo_off, int use_lo, int HW,
                   const float* __restrict__ hw, const float* __restrict__ hb, double* __restrict__ stats_partial) {
  __shared__ float sw[K * 32];
  __shared__ float sb[K];
  for (int i = threadIdx.x; i < K * 32; i += blockDim.x) sw[i] = hw[i];
  if (threadIdx.x < K) sb[threadIdx.x] = hb[threadIdx.x];
  __syncthreads();
  const int img = blockIdx.y;
  double acc[3 * K];
#pragma unroll
  for (int i = 0; i < 3 * K; ++i) acc[i] = 0.0;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    float f[32], p[K];
    int y;
    badge_pixel<K>(feat + ((size_t)img * HW + px) * (size_t)pitch, hi_off, lo_off, use_lo, sw, sb, f, p, y);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      acc[K + k] += (double)p[k];
      if (k == y) {
        acc[k] += (double)p[k];
        acc[2 * K + k] += 1.0;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 3 * K; ++i) {
    const double r = block_sum(acc[i]);
    if (threadIdx.x == 0) stats_partial[((size_t)img * gridDim.x + blockIdx.x) * (3 * K) + i] = r;
    __syncthreads();
  }
}

// <<<ceil(n/128), 128>>>  q coefficients per image: coef [n][2K] = (a_c, b_c) with q_c = a_c*y_c + b_c
template <int K>
__global__ void badge_coef_kernel(const double* __restrict__ stats_partial, int blocks, float smooth, int do_bg,
                                  float* __restrict__ coef, int n) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  if (img >= n) return;
  const int C = do_bg ? K : K - 1;
  for (int k = 0; k < K; ++k) {
    double I = 0.0, S = 0.0, T = 0.0;
    for (int b = 0; b < blocks; ++b) {
      const double* s = stats_partial + ((size_t)img * blocks + b) * (3 * K);
      I += s[k];
      S += s[K + k];
      T += s[2 * K + k];
    }
    const double U = S + T + (double)smooth;
    const bool used = do_bg || k > 0;
    coef[(size_t)img * 2 * K + k] = used ? (float)(-2.0 / (U * C)) : 0.f;
    coef[(size_t)img * 2 * K + K + k] = used ? (float)((2.0 * I + (double)smooth) / (U * U * C)) : 0.f;
  }
}

// grid (kBadgeBlocks, n); grad_partial [n][kBadgeBlocks][K*32] doubles
template <int K>
__global__ void __launch_bounds__(256)
badge_grad_kernel(const __nv_bfloat16* __restrict__ feat, int pitch, int hi_off, int lo_off, int use_lo, int HW,
                  const float* __restrict__ hw, const float* __restrict__ hb, const float* __restrict__ coef,
                  double* __restrict__ grad_partial) {
  __shared__ float sw[K * 32];
  __shared__ float sb[K];
  __shared__ float red[8][K * 32];
  for (int i = threadIdx.x; i < K * 32; i += blockDim.x) sw[i] = hw[i];
  if (threadIdx.x < K) sb[threadIdx.x] = hb[threadIdx.x];
  __syncthreads();
  const int img = blockIdx.y;
  float qa[K], qb[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    qa[k] = coef[(size_t)img * 2 * K + k];
    qb[k] = coef[(size_t)img * 2 * K + K + k];
  }
  const float invP = 1.0f / (float)HW;
  float G[K][32];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int c = 0; c < 32; ++c) G[k][c] = 0.f;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    float f[32], p[K];
    int y;
    badge_pixel<K>(feat + ((size_t)img * HW + px) * (size_t)pitch, hi_off, lo_off, use_lo, sw, sb, f, p, y);
    float q[K], pq = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      q[k] = (k == y ? qa[k] : 0.f) + qb[k];
      pq = fmaf(p[k], q[k], pq);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float g = (p[k] - (k == y ? 1.f : 0.f)) * invP + p[k] * (q[k] - pq);
#pragma unroll
      for (int c = 0; c < 32; ++c) G[k][c] = fmaf(g, f[c], G[k][c]);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      float v = G[k][c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) red[warp][k * 32 + c] = v;
    }
  __syncthreads();
  if (threadIdx.x < K * 32) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += (double)red[w][threadIdx.x];
    grad_partial[((size_t)img * gridDim.x + blockIdx.x) * (K * 32) + threadIdx.x] = s;
  }
}

// out [n][out_stride] fp32: K*32 gradient entries, zero-padded to out_stride
__global__ void badge_reduce_kernel(const double* __restrict__ grad_partial, int blocks, int kd, int out_stride,
                                    float* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * out_stride) return;
  const int img = i / out_stride, j = i % out_stride;
  double s = 0.0;
  if (j < kd)
    for (int b = 0; b < blocks; ++b) s += grad_partial[((size_t)img * blocks + b) * kd + j];
  out[i] = (float)s;
}

}  // namespace b200al
