// Per-image uncertainty scores without materialising the probability map.
//   score_entropy    = mean_px mean_k ( -p_k * log2(p_k + smooth) )      entropy_selector.py:45-49
//   score_margin     = mean_px -( p_top1 - p_top2 )                      margin_selector.py:45-48
//   score_confidence = mean_px -max_k p_k                                confidence_selector.py:45-47
// with p = softmax over classes, optionally averaged over T MC-dropout repeats
// before the score formula.  Two front ends share the per-pixel math:
//   * uncertainty_logits_kernel: fp32 logits [n, T, K, HW] in HBM (kernel 3 of
//     the north star, the HBM-roofline kernel): float4 loads, one pass;
//   * head_uncertainty_kernel: fused with the UNet's 1x1 seg_output conv, reads
//     the last decoder feature map (32 ch, hi/lo bf16) so logits never reach HBM.
// Reductions are fixed-order (thread -> warp shuffle -> block -> per-image
// partials summed in index order in fp64): run-to-run bit-identical.
#pragma once

#include "common.cuh"

namespace b200al {

enum { MODE_ENTROPY = 0, MODE_MARGIN = 1, MODE_CONFIDENCE = 2, MODE_DICE = 3 };
constexpr int kDiceBlocks = 32;  // partial blocks per image of the dice-uncertainty reduction
constexpr int kMaxClasses = 8;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// pbar += softmax(l).  exp2/rcp/log2 run on the SFU (ex2.approx / rcp.approx / lg2.approx, <= 2 ulp):
// with IEEE expf and three divisions per pixel the reduction was instruction-bound at 0.39 of the HBM
// roofline; the approximations are four orders of magnitude inside the 1e-3 score tolerance.
template <int K>
__device__ __forceinline__ void softmax_accum(const float (&l)[K], float (&pbar)[K]) {
  float m = l[0];
#pragma unroll
  for (int k = 1; k < K; ++k) m = fmaxf(m, l[k]);
  constexpr float kLog2e = 1.4426950408889634f;
  const float mc = m * kLog2e;
  float e[K], sum = 0.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    e[k] = ex2_approx(fmaf(l[k], kLog2e, -mc));
    sum += e[k];
  }
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(sum));
#pragma unroll
  for (int k = 0; k < K; ++k) pbar[k] = fmaf(e[k], r, pbar[k]);
}

template <int K>
__device__ __forceinline__ float score_term(const float (&p)[K], int mode, float smooth) {
  if (mode == MODE_ENTROPY) {
    float h = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) h = fmaf(-p[k], __log2f(p[k] + smooth), h);
    return h / (float)K;
  }
  float t1 = -1.f, t2 = -1.f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (p[k] > t1) { t2 = t1; t1 = p[k]; }
    else if (p[k] > t2) t2 = p[k];
  }
  return mode == MODE_MARGIN ? -(t1 - t2) : -t1;
}

// block-wide fixed-order sum in fp64; result valid in thread 0
__device__ __forceinline__ double block_sum(double v) {
  __shared__ double warp_part[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) warp_part[w] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < (int)((blockDim.x + 31) >> 5); ++i) r += warp_part[i];
  return r;
}

// logits [n, T, K, HW] fp32; grid (blocks_per_image, n); partial [n][gridDim.x]
template <int K>
__global__ void __launch_bounds__(256)
uncertainty_logits_kernel(const float* __restrict__ logits, int T, int HW, int mode, float smooth,
                          double* __restrict__ partial) {
  const int img = blockIdx.y;
  const float* base = logits + (size_t)img * T * K * HW;
  double acc = 0.0;
  const float inv_T = 1.0f / (float)T;
  for (int p4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; p4 < HW; p4 += gridDim.x * blockDim.x * 4) {
    float pbar[4][K];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < K; ++k) pbar[j][k] = 0.f;
    for (int t = 0; t < T; ++t) {
      float4 v[K];
#pragma unroll
      for (int k = 0; k < K; ++k)
        v[k] = __ldg(reinterpret_cast<const float4*>(base + ((size_t)t * K + k) * HW + p4));
      float l[K];
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].x;
      softmax_accum<K>(l, pbar[0]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].y;
      softmax_accum<K>(l, pbar[1]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].z;
      softmax_accum<K>(l, pbar[2]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].w;
      softmax_accum<K>(l, pbar[3]);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (T > 1) {
#pragma unroll
        for (int k = 0; k < K; ++k) pbar[j][k] *= inv_T;
      }
      s += score_term<K>(pbar[j], mode, smooth);
    }
    acc += (double)s;
  }
  const double r = block_sum(acc);
  if (threadIdx.x == 0) partial[(size_t)img * gridDim.x + blockIdx.x] = r;
}

// score[img] = (sum of partials in index order) / HW
__global__ void uncertainty_finalize_kernel(const double* __restrict__ partial, int nblk, int HW,
                                            float* __restrict__ scores, int n) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  if (img >= n) return;
  double s = 0.0;
  for (int i = 0; i < nblk; ++i) s += partial[(size_t)img * nblk + i];
  scores[img] = (float)(s / (double)HW);
}

// Fused 1x1 head + softmax + score.  feat: [S = n*T, HW, pitch] bf16 hi/lo, 32 ch.
// head weight [K][32], bias [K] fp32.  logits_out nullable: [S, K, HW] fp32.
template <int K>
__global__ void __launch_bounds__(256)
head_uncertainty_kernel(const __nv_bfloat16* __restrict__ feat, int pitch, int hi_off, int lo_off, int use_lo,
                        int HW, int T, const float* __restrict__ hw, const float* __restrict__ hb, int mode,
                        float smooth, float* __restrict__ logits_out, double* __restrict__ partial) {
  __shared__ float sw[K * 32];
  __shared__ float sb[K];
  for (int i = threadIdx.x; i < K * 32; i += blockDim.x) sw[i] = hw[i];
  if (threadIdx.x < K) sb[threadIdx.x] = hb[threadIdx.x];
  __syncthreads();
  const int img = blockIdx.y;
  double acc = 0.0;
  const float inv_T = 1.0f / (float)T;
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < HW; px += gridDim.x * blockDim.x) {
    float pbar[K];
#pragma unroll
    for (int k = 0; k < K; ++k) pbar[k] = 0.f;
    for (int t = 0; t < T; ++t) {
      const size_t s = (size_t)img * T + t;
      const __nv_bfloat16* src = feat + (s * HW + px) * (size_t)pitch;
      float f[32];
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
      float l[K];
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float a = sb[k];
#pragma unroll
        for (int c = 0; c < 32; ++c) a = fmaf(f[c], sw[k * 32 + c], a);
        l[k] = a;
        if (logits_out) logits_out[(s * K + k) * (size_t)HW + px] = a;
      }
      softmax_accum<K>(l, pbar);
    }
    if (T > 1) {
#pragma unroll
      for (int k = 0; k < K; ++k) pbar[k] *= inv_T;
    }
    acc += (double)score_term<K>(pbar, mode, smooth);
  }
  const double r = block_sum(acc);
  if (threadIdx.x == 0) partial[(size_t)img * gridDim.x + blockIdx.x] = r;
}

// ---------------------------------------------------------------------------------------------------------------
// Dice-uncertainty (north-star kernel 3, third mode): score = DiceLoss(pbar, one_hot(argmax_k pbar)) with
// the reference's DiceLoss semantics (src/losses/dice_loss.py:32-72: do_bg = True, batch = False, squared = False):
//   I_c = sum_px pbar_c * y_c,  P_c = sum_px pbar_c,  Y_c = sum_px y_c,
//   score = mean_c ( 1 - (2 I_c + smooth) / (P_c + Y_c + smooth) ).
// One pass over the logits like the other modes, but 3K running sums per image instead of one; the non-linear part
// runs in the finalize kernel on the per-image totals.  partial: [n][gridDim.x][3K] doubles.
template <int K>
__global__ void __launch_bounds__(256)
dice_logits_kernel(const float* __restrict__ logits, int T, int HW, double* __restrict__ partial) {
  const int img = blockIdx.y;
  const float* base = logits + (size_t)img * T * K * HW;
  float accI[K], accP[K], accY[K];
#pragma unroll
  for (int k = 0; k < K; ++k) accI[k] = accP[k] = accY[k] = 0.f;
  const float inv_T = 1.0f / (float)T;
  for (int p4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4; p4 < HW; p4 += gridDim.x * blockDim.x * 4) {
    float pbar[4][K];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < K; ++k) pbar[j][k] = 0.f;
    for (int t = 0; t < T; ++t) {
      float4 v[K];
#pragma unroll
      for (int k = 0; k < K; ++k)
        v[k] = __ldg(reinterpret_cast<const float4*>(base + ((size_t)t * K + k) * HW + p4));
      float l[K];
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].x;
      softmax_accum<K>(l, pbar[0]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].y;
      softmax_accum<K>(l, pbar[1]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].z;
      softmax_accum<K>(l, pbar[2]);
#pragma unroll
      for (int k = 0; k < K; ++k) l[k] = v[k].w;
      softmax_accum<K>(l, pbar[3]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int am = 0;
      float best = pbar[j][0];
#pragma unroll
      for (int k = 1; k < K; ++k)
        if (pbar[j][k] > best) { best = pbar[j][k]; am = k; }  // first maximum, like torch.argmax
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float p = T > 1 ? pbar[j][k] * inv_T : pbar[j][k];
        accP[k] += p;
        if (k == am) { accI[k] += p; accY[k] += 1.f; }
      }
    }
  }
  // a thread adds at most HW / (4 * 256) * 4 terms in fp32 (<= 256 at 256^2 with one block); the rest is fp64
  double* out = partial + ((size_t)img * gridDim.x + blockIdx.x) * 3 * K;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double r = block_sum((double)accI[k]);
    if (threadIdx.x == 0) out[k] = r;
    __syncthreads();
    r = block_sum((double)accP[k]);
    if (threadIdx.x == 0) out[K + k] = r;
    __syncthreads();
    r = block_sum((double)accY[k]);
    if (threadIdx.x == 0) out[2 * K + k] = r;
    __syncthreads();
  }
}

template <int K>
__global__ void dice_finalize_kernel(const double* __restrict__ partial, int nblk, float smooth,
                                     float* __restrict__ scores, int n) {
  const int img = blockIdx.x * blockDim.x + threadIdx.x;
  if (img >= n) return;
  double loss = 0.0;
  for (int k = 0; k < K; ++k) {
    double I = 0.0, P = 0.0, Y = 0.0;
    for (int b = 0; b < nblk; ++b) {
      const double* p = partial + ((size_t)img * nblk + b) * 3 * K;
      I += p[k]; P += p[K + k]; Y += p[2 * K + k];
    }
    loss += 1.0 - (2.0 * I + (double)smooth) / (P + Y + (double)smooth);
  }
  scores[img] = (float)(loss / (double)K);
}

}  // namespace b200al
