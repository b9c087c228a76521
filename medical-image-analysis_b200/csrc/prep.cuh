// Pool-image input pipeline on the device: grayscale uint8 images of ragged sizes -> the fp32 [H, W] image the model
// sees, i.e. the reference's dataset-side `F.to_tensor` -> `JointResize` -> `ZScoreNormalize` chain
// (src/datasets/fugc/fugc_dataset.py:140-160, src/transforms/joint_transform.py:20-25,
// src/transforms/normalization.py:13-21).
//
// JointResize on a float tensor is torch's antialiased bilinear interpolate (align_corners=False): a separable
// triangle filter whose support grows with the down-scale factor.  Per output index i of an axis (fp32 arithmetic):
//   scale = in/out, support = max(scale, 1), center = scale*(i+0.5), invscale = 1/max(scale, 1)
//   lo = max(0, int(center-support+0.5)), n = min(in, int(center+support+0.5)) - lo
//   w_j = max(0, 1-|(j+lo-center+0.5)*invscale|) / sum_j(...)
// Horizontal pass first, then vertical, sequential fp32 accumulation in tap order (oracle/prep_ref.py restates the
// same and is pinned against torchvision).
//
// One CTA = one band of kPrepBand output rows of one image, one thread per output column.  Each source row that the
// band touches is filtered horizontally once (registers) and scattered into the <= kPrepBand vertical accumulators
// it contributes to, so no intermediate image is written.  The source bytes are read straight from the ragged pool
// cache (L2-friendly: neighbouring bands share rows).  Per-band sum / sum of squares go to a partials array so the
// z-score statistics are reduced in a fixed order (bit-reproducible selection).
#pragma once
#include <cstdint>

#include "common.cuh"

namespace b200al {

constexpr int kPrepBand = 8;
constexpr int kPrepMaxTaps = 35;  // 2*ceil(scale)+1 with scale <= 17

struct PrepTap {
  int lo, n;
};

// taps of output index i along an axis of `in` source pixels; weights to w[0..n) with stride `ws`
__device__ __forceinline__ PrepTap prep_taps(int i, int in, float scale, int max_taps, float* w, int ws) {
  const float support = scale >= 1.f ? scale : 1.f;
  const float invscale = scale >= 1.f ? __fdiv_rn(1.f, scale) : 1.f;
  const float center = __fmul_rn(scale, (float)i + 0.5f);
  int lo = (int)(__fadd_rn(__fsub_rn(center, support), 0.5f));
  if (lo < 0) lo = 0;
  int hi = (int)(__fadd_rn(__fadd_rn(center, support), 0.5f));
  if (hi > in) hi = in;
  int n = hi - lo;
  n = n < 0 ? 0 : (n > max_taps ? max_taps : n);
  float tot = 0.f;
  for (int j = 0; j < n; ++j) {
    const float x = __fmul_rn(__fadd_rn(__fsub_rn(__fadd_rn((float)j, (float)lo), center), 0.5f), invscale);
    float v = 1.f - fabsf(x);
    v = v > 0.f ? v : 0.f;
    w[j * ws] = v;
    tot = __fadd_rn(tot, v);
  }
  if (tot != 0.f) {
    const float inv = __fdiv_rn(1.f, tot);
    for (int j = 0; j < n; ++j) w[j * ws] = __fmul_rn(w[j * ws], inv);
  }
  return {lo, n};
}

// grid (ceil(Ho / kPrepBand), n_images), block Wo threads (Wo <= 1024); dynamic smem = (max_taps*Wo + kPrepBand*max_taps) floats
__global__ void prep_resize_kernel(const uint8_t* __restrict__ raw, const int64_t* __restrict__ offsets,
                                   const int32_t* __restrict__ hw, int Ho, int Wo, int max_taps,
                                   float* __restrict__ out, double* __restrict__ partials) {
  extern __shared__ float prep_sm[];
  float* wx = prep_sm;                   // [max_taps][Wo]
  float* wy = prep_sm + max_taps * Wo;   // [kPrepBand][max_taps]
  __shared__ PrepTap ty[kPrepBand];
  __shared__ double red[2][32];
  const int img = blockIdx.y, band = blockIdx.x;
  const int Hs = hw[2 * img], Ws = hw[2 * img + 1];
  const uint8_t* src = raw + offsets[img];
  const int x = threadIdx.x;
  const int y0 = band * kPrepBand;
  const PrepTap tx = prep_taps(x, Ws, (float)Ws / (float)Wo, max_taps, wx + x, Wo);
  if (x < kPrepBand) {
    PrepTap t = {0, 0};
    if (y0 + x < Ho) t = prep_taps(y0 + x, Hs, (float)Hs / (float)Ho, max_taps, wy + x * max_taps, 1);
    ty[x] = t;
  }
  __syncthreads();
  int r0 = Hs, r1 = 0;
#pragma unroll
  for (int k = 0; k < kPrepBand; ++k)
    if (ty[k].n > 0) {
      r0 = min(r0, ty[k].lo);
      r1 = max(r1, ty[k].lo + ty[k].n);
    }
  float acc[kPrepBand];
#pragma unroll
  for (int k = 0; k < kPrepBand; ++k) acc[k] = 0.f;
  for (int r = r0; r < r1; ++r) {
    const uint8_t* row = src + (size_t)r * Ws + tx.lo;
    float h = 0.f;
    for (int j = 0; j < tx.n; ++j) h = __fmaf_rn(__fdiv_rn((float)row[j], 255.0f), wx[j * Wo + x], h);  // to_tensor: u8/255
#pragma unroll
    for (int k = 0; k < kPrepBand; ++k) {
      const int j = r - ty[k].lo;
      if (j >= 0 && j < ty[k].n) acc[k] = __fmaf_rn(h, wy[k * max_taps + j], acc[k]);
    }
  }
  double s = 0.0, ss = 0.0;
#pragma unroll
  for (int k = 0; k < kPrepBand; ++k)
    if (y0 + k < Ho) {
      out[((size_t)img * Ho + y0 + k) * Wo + x] = acc[k];
      s += (double)acc[k];
      ss += (double)acc[k] * (double)acc[k];
    }
  if (partials) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      ss += __shfl_xor_sync(0xffffffffu, ss, o);
    }
    if ((x & 31) == 0) {
      red[0][x >> 5] = s;
      red[1][x >> 5] = ss;
    }
    __syncthreads();
    if (x == 0) {
      double a = 0.0, b = 0.0;
      for (int i = 0; i < (int)(blockDim.x + 31) / 32; ++i) {
        a += red[0][i];
        b += red[1][i];
      }
      partials[((size_t)img * gridDim.x + band) * 2] = a;
      partials[((size_t)img * gridDim.x + band) * 2 + 1] = b;
    }
  }
}

// ZScoreNormalize statistics (normalization.py:17-21): mean and UNBIASED std over the H*W pixels, std clipped at 1e-8.
// norm[img] = (mean, std); the first conv's loader applies (x - mean) / std.
__global__ void prep_stats_kernel(const double* __restrict__ partials, int bands, int count, float* __restrict__ norm,
                                  int n) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  if (img >= n) return;
  double s = 0.0, ss = 0.0;
  for (int b = 0; b < bands; ++b) {
    s += partials[((size_t)img * bands + b) * 2];
    ss += partials[((size_t)img * bands + b) * 2 + 1];
  }
  const double mean = s / count;
  double var = (ss - s * mean) / (count - 1);
  if (var < 0.0) var = 0.0;
  const float sd = (float)sqrt(var);
  norm[2 * img] = (float)mean;
  norm[2 * img + 1] = sd > 1e-8f ? sd : 1e-8f;
}

}  // namespace b200al
