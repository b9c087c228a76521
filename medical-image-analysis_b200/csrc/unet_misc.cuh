// UNet kernels that are not GEMM-shaped: weight packing / BN folding, the
// Cin=1 first conv (K=9, CUDA cores, fp32), MC-dropout mask generation, the
// bottleneck average pool, and a plain SIMT convolution used only by tests to
// cross-check the tcgen05 path on the same packed operands.
#pragma once

#include "common.cuh"

namespace b200al {

// Conv2d weight [Cout][Cin][3][3] fp32 -> packed [plane][tap][Cout][Cin] bf16 (hi, lo).
__global__ void pack_conv3x3_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int cout, int cin) {
  const int total = cout * cin * 9;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int tap = i % 9;
    const int ci = (i / 9) % cin;
    const int co = i / (9 * cin);
    __nv_bfloat16 hi, lo;
    split_bf16(w[i], hi, lo);
    const size_t o = ((size_t)tap * cout + co) * cin + ci;
    wp[o] = hi;
    wp[(size_t)9 * cout * cin + o] = lo;
  }
}

// Conv2d weight [Cout][32][3][3] fp32 -> [tap][Cout][hi 32 | lo 32] bf16: 128-byte rows for the
// packed variant of the column-folded kernel (conv_fold.cuh).
__global__ void pack_conv3x3_rowpair_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int cout) {
  const int total = cout * 32 * 9;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int tap = i % 9;
    const int ci = (i / 9) % 32;
    const int co = i / (9 * 32);
    __nv_bfloat16 hi, lo;
    split_bf16(w[i], hi, lo);
    const size_t o = ((size_t)tap * cout + co) * 64 + ci;
    wp[o] = hi;
    wp[o + 32] = lo;
  }
}

// ConvTranspose2d weight [Cin][Cout][2][2] -> packed [plane][quad*Cout + co][Cin], quad = dy*2+dx.
__global__ void pack_convT_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp, int cin, int cout) {
  const int total = cin * cout * 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int quad = i % 4;
    const int co = (i / 4) % cout;
    const int ci = i / (4 * cout);
    __nv_bfloat16 hi, lo;
    split_bf16(w[i], hi, lo);
    const size_t o = ((size_t)quad * cout + co) * cin + ci;
    wp[o] = hi;
    wp[(size_t)4 * cout * cin + o] = lo;
  }
}

// Single-plane fp16 packings of the same two layouts (plane 0 only) for the fp16 layers of the mixed-precision
// scoring path (conv_tc.cuh, FMT >= 2).
__global__ void pack_conv3x3_f16_kernel(const float* __restrict__ w, __half* __restrict__ wp, int cout, int cin) {
  const int total = cout * cin * 9;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int tap = i % 9;
    const int ci = (i / 9) % cin;
    const int co = i / (9 * cin);
    wp[((size_t)tap * cout + co) * cin + ci] = __float2half_rn(w[i]);
  }
}
__global__ void pack_convT_f16_kernel(const float* __restrict__ w, __half* __restrict__ wp, int cin, int cout) {
  const int total = cin * cout * 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int quad = i % 4;
    const int co = (i / 4) % cout;
    const int ci = i / (4 * cout);
    wp[((size_t)quad * cout + co) * cin + ci] = __float2half_rn(w[i]);
  }
}

// BatchNorm (eval) folded to y = (acc + bias) * scale + shift (blocks.py:98;
// Dropout2d sits between conv and BN, so the MC mask multiplies `scale`).
__global__ void fold_bn_kernel(const float* gamma, const float* beta, const float* mean, const float* var, float eps,
                               float* scale, float* shift, int c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < c) {
    const float s = gamma[i] / sqrtf(var[i] + eps);
    scale[i] = s;
    shift[i] = beta[i] - mean[i] * s;
  }
}

__global__ void fill_kernel(float* p, float v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// mask[s][ch] for s = b*T + t: 0 or 1/(1-p); keep iff hash24 >= thr.
__global__ void mc_mask_kernel(float* __restrict__ mask, int n_samples, int T, int n_ch, uint64_t seed,
                               long long base_pos, uint32_t thr, float keep_scale) {
  const long long total = (long long)n_samples * n_ch;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(i % n_ch);
    const int s = (int)(i / n_ch);
    const int b = s / T, t = s - b * T;
    mask[i] = mc_hash24(seed, (uint64_t)(base_pos + b), (uint32_t)t, (uint32_t)ch) >= thr ? keep_scale : 0.f;
  }
}

// First conv: 1 -> 32 channels, 3x3, pad 1, fp32 math on CUDA cores (K = 9).
// One thread per output pixel of one IMAGE: the 32 pre-activation sums are computed
// once and reused by the T MC-dropout repeats (samples img*T .. img*T+T-1), which
// differ only in the per-sample channel mask folded into the BN scale.  The filter
// and the folded BN constants travel as a kernel parameter, so every FMA takes its
// weight from the constant bank instead of a shared-memory load.
struct FirstConvConsts {
  float w[32 * 9];
  float bias[32], scale[32], shift[32];
};

// two fp32 values -> packed bf16x2 (a in the low half), round to nearest even: one F2FP instead of two F2F + a PRMT
__device__ __forceinline__ uint32_t cvt_bf16x2_rn(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}

// UNIT_SLOPE: 0 <= slope <= 1 (the reference's 0.01, blocks.py:100), where leaky(y) == max(y, slope * y)
template <typename ImgT, bool UNIT_SLOPE>
__global__ void __launch_bounds__(256, 3)
conv_first_kernel(const ImgT* __restrict__ img, const float* __restrict__ norm, int T, int H, int W,
                  const __grid_constant__ FirstConvConsts k,
                  const float* __restrict__ mask, int mask_stride, float slope, __nv_bfloat16* __restrict__ out,
                  int out_pitch, int hi_off, int lo_off, int write_lo, int rows) {
  extern __shared__ __align__(16) float se[];  // [T][32] scale, then [T][32] shift
  float* ses = se;
  float* seh = se + T * 32;
  const int im_idx = blockIdx.z;
  // launched with 256 threads (run_first_conv)
  for (int i = threadIdx.x; i < T * 32; i += 256) {
    const int t = i >> 5, c = i & 31;
    float sc = k.scale[c];
    if (mask) sc *= mask[(size_t)(im_idx * T + t) * mask_stride + c];
    ses[i] = sc;
    seh[i] = k.bias[c] * sc + k.shift[c];
  }
  // ZScoreNormalize fused into the loader (normalization.py:21): (x - mean) / std on real pixels, padding stays 0
  const float nm = norm ? norm[2 * im_idx] : 0.f;
  const float nsd = norm ? norm[2 * im_idx + 1] : 1.f;
  // 8-bit pixels: torchvision to_tensor (x / 255, an IEEE division) and the z-score division, once per byte value
  // and block instead of nine times per pixel
  __shared__ float lut[256];
  __shared__ __align__(16) float wsm[288];
  for (int i = threadIdx.x; i < 288; i += 256) wsm[i] = k.w[i];
  if (sizeof(ImgT) == 1) {
    float t = (float)threadIdx.x / 255.0f;
    if (norm) t = __fdiv_rn(t - nm, nsd);
    lut[threadIdx.x] = t;
  }
  __syncthreads();
  // every store instruction of a warp must cover contiguous bytes (a lane-strided 16-byte store costs
  // one L1 wavefront per lane): the 32 pixels x 128 B of a warp are transposed through shared memory
  __shared__ uint4 stg[8][32 * 8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int x0w = x - lane;
  const ImgT* im = img + (size_t)im_idx * H * W;
  // A block walks `rows` consecutive image rows: the 3x3 window slides down (three new taps per row instead of
  // nine), the lookup table / weights / scales are set up once, and the next row's bytes are in flight while this
  // row is computed and stored.  Taps are clamped unconditional loads, padding is selected to 0 afterwards.
  int xc[3];
  bool cok[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int xx = x + c - 1;
    cok[c] = (unsigned)xx < (unsigned)W;
    xc[c] = min(max(xx, 0), W - 1);
  }
  auto fetch = [&](int yy, ImgT* raw) {
    const ImgT* row = im + (size_t)min(max(yy, 0), H - 1) * W;
#pragma unroll
    for (int c = 0; c < 3; ++c) raw[c] = row[xc[c]];
  };
  auto taps = [&](int yy, const ImgT* raw, float* dst) {
    const bool rok = (unsigned)yy < (unsigned)H;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float t;
      if (sizeof(ImgT) == 1) {
        t = lut[(int)raw[c]];
      } else {
        t = (float)raw[c];
        if (norm) t = __fdiv_rn(t - nm, nsd);
      }
      dst[c] = (rok && cok[c]) ? t : 0.f;
    }
  };
  const int y0 = blockIdx.y * rows;
  const int y1 = min(y0 + rows, H);
  float v[9];
  ImgT raw[3];
  fetch(y0 - 1, raw);
  taps(y0 - 1, raw, v + 3);
  fetch(y0, raw);
  taps(y0, raw, v + 6);
  fetch(y0 + 1, raw);
  uint4* ws = stg[warp];
  for (int y = y0; y < y1; ++y) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    v[c] = v[3 + c];
    v[3 + c] = v[6 + c];
  }
  taps(y + 1, raw, v + 6);
  fetch(y + 2, raw);  // for the next row
  // weights from shared memory, four per broadcast load: as kernel-parameter operands the 288 weights went through
  // uniform registers and cost about one extra instruction per FFMA in moves and refills
  float acc[32];
#pragma unroll
  for (int co = 0; co < 32; ++co) acc[co] = 0.f;
  const float4* w4 = reinterpret_cast<const float4*>(wsm);
#pragma unroll
  for (int q = 0; q < 72; ++q) {
    const float4 w = w4[q];
    const float we[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int e = 4 * q + j;  // = co * 9 + tap, taps of a channel in ascending order as before
      acc[e / 9] = fmaf(v[e % 9], we[j], acc[e / 9]);
    }
  }
  for (int t = 0; t < T; ++t) {
    const float* sc = ses + t * 32;
    const float* sh = seh + t * 32;
    __syncwarp();  // the previous repeat's reads of ws are done
#pragma unroll
    for (int g = 0; g < 4; ++g) {  // 8 channels at a time: one 16-byte hi word and one lo word per pixel
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int co = g * 8 + q * 4;
        const float4 s4 = *reinterpret_cast<const float4*>(sc + co);
        const float4 h4 = *reinterpret_cast<const float4*>(sh + co);
        float a[4] = {fmaf(acc[co], s4.x, h4.x), fmaf(acc[co + 1], s4.y, h4.y), fmaf(acc[co + 2], s4.z, h4.z),
                      fmaf(acc[co + 3], s4.w, h4.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) a[j] = UNIT_SLOPE ? fmaxf(a[j], a[j] * slope) : leaky(a[j], slope);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const uint32_t h = cvt_bf16x2_rn(a[2 * j], a[2 * j + 1]);
          hi[q * 2 + j] = h;
          lo[q * 2 + j] = cvt_bf16x2_rn(a[2 * j] - bf16lo_to_float(h), a[2 * j + 1] - bf16hi_to_float(h));
        }
      }
      ws[lane * 8 + (g ^ (lane & 7))] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      ws[lane * 8 + ((4 + g) ^ (lane & 7))] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
    __syncwarp();
    // lane -> (pixel lane >> 3 of each group of 4, 16-byte column c of the pixel's 128 B); 8 stores, 4 pixels apart
    const int c = lane & 7;
    __nv_bfloat16* d = out + ((size_t)((size_t)(im_idx * T + t) * H + y) * W + x0w + (lane >> 3)) * out_pitch +
                       (c < 4 ? hi_off + c * 8 : lo_off + (c - 4) * 8);
    const bool col_on = c < 4 || write_lo;
    const uint4* wsr = ws + (lane >> 3) * 8;
    if (x0w + 32 <= W) {  // whole warp inside the row (always, when W is a multiple of 32)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int px = i * 4 + (lane >> 3);
        const uint4 val = wsr[i * 32 + (c ^ (px & 7))];
        if (col_on) *reinterpret_cast<uint4*>(d + (size_t)i * 4 * out_pitch) = val;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int px = i * 4 + (lane >> 3);
        const uint4 val = wsr[i * 32 + (c ^ (px & 7))];
        if (x0w + px < W && col_on) *reinterpret_cast<uint4*>(d + (size_t)i * 4 * out_pitch) = val;
      }
    }
  }
  }  // rows
}

// Bottleneck global average pool (unet.py:92): [S, HW, pitch] hi/lo -> [S, C] fp32.
__global__ void avgpool_kernel(const __nv_bfloat16* __restrict__ act, int hw, int pitch, int hi_off, int lo_off,
                               int use_lo, int c_total, float* __restrict__ emb) {
  const int s = blockIdx.x;
  for (int c = threadIdx.x; c < c_total; c += blockDim.x) {
    const __nv_bfloat16* base = act + (size_t)s * hw * pitch + c;
    float acc = 0.f;
    int i = 0;
    // same pixel-order sum as the plain loop below (bit-identical), with 16 pixels' loads in flight at a time: the
    // kernel is bound by load latency (one block per sample, 256 dependent steps), not by bytes
    for (; i + 16 <= hw; i += 16) {
      float vh[16], vl[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        vh[j] = __bfloat162float(base[(size_t)(i + j) * pitch + hi_off]);
        vl[j] = use_lo ? __bfloat162float(base[(size_t)(i + j) * pitch + lo_off]) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) acc += use_lo ? vh[j] + vl[j] : vh[j];
    }
    for (; i < hw; ++i) {
      float v = __bfloat162float(base[(size_t)i * pitch + hi_off]);
      if (use_lo) v += __bfloat162float(base[(size_t)i * pitch + lo_off]);
      acc += v;
    }
    emb[(size_t)s * c_total + c] = acc / (float)hw;
  }
}

// TEST-ONLY: direct convolution on CUDA cores over the same NHWC hi/lo
// activations and packed weights the tcgen05 kernel consumes.  fp32 FMA.
// One thread per (output pixel, cout).  mode 0: 3x3 conv (stride 1|2, pad 1);
// mode 1: 2x2 stride-2 transposed conv.
struct ConvSimtParams {
  const __nv_bfloat16* in;
  long long in_pitch;
  int in_lo_off, in_H, in_W, cin;
  const __nv_bfloat16* wp;  // packed as for the tcgen05 kernel
  int cout, stride, mode, use_lo;
  const float *scale, *bias, *shift, *mask;
  int mask_stride;
  float slope;
  __nv_bfloat16* out;
  long long out_pitch;
  int out_hi_off, out_lo_off, out_H, out_W, write_lo;
};

__global__ void conv_simt_kernel(const ConvSimtParams p, int n_samples) {
  const long long total = (long long)n_samples * p.out_H * p.out_W * p.cout;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int co = (int)(i % p.cout);
    long long pix = i / p.cout;
    const int ox = (int)(pix % p.out_W);
    pix /= p.out_W;
    const int oy = (int)(pix % p.out_H);
    const int s = (int)(pix / p.out_H);
    float acc = 0.f;
    const int taps = p.mode == 0 ? 9 : 4;
    const size_t plane = (size_t)taps * p.cout * p.cin;
    if (p.mode == 0) {
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
          const int iy = oy * p.stride + r - 1, ix = ox * p.stride + c - 1;
          if (iy < 0 || iy >= p.in_H || ix < 0 || ix >= p.in_W) continue;
          const __nv_bfloat16* a = p.in + ((size_t)((size_t)s * p.in_H + iy) * p.in_W + ix) * p.in_pitch;
          const __nv_bfloat16* w = p.wp + ((size_t)(r * 3 + c) * p.cout + co) * p.cin;
          for (int ci = 0; ci < p.cin; ++ci) {
            float av = __bfloat162float(a[ci]);
            float wv = __bfloat162float(w[ci]);
            if (p.use_lo) {
              av += __bfloat162float(a[ci + p.in_lo_off]);
              wv += __bfloat162float(w[plane + ci]);
            }
            acc = fmaf(av, wv, acc);
          }
        }
    } else {
      const int iy = oy >> 1, ix = ox >> 1, quad = (oy & 1) * 2 + (ox & 1);
      const __nv_bfloat16* a = p.in + ((size_t)((size_t)s * p.in_H + iy) * p.in_W + ix) * p.in_pitch;
      const __nv_bfloat16* w = p.wp + ((size_t)quad * p.cout + co) * p.cin;
      for (int ci = 0; ci < p.cin; ++ci) {
        float av = __bfloat162float(a[ci]);
        float wv = __bfloat162float(w[ci]);
        if (p.use_lo) {
          av += __bfloat162float(a[ci + p.in_lo_off]);
          wv += __bfloat162float(w[plane + ci]);
        }
        acc = fmaf(av, wv, acc);
      }
    }
    float sc = p.scale[co];
    if (p.mask) sc *= p.mask[(size_t)s * p.mask_stride + co];
    const float y = leaky(acc * sc + (p.bias[co] * sc + p.shift[co]), p.slope);
    __nv_bfloat16 hi, lo;
    split_bf16(y, hi, lo);
    __nv_bfloat16* dst = p.out + ((size_t)((size_t)s * p.out_H + oy) * p.out_W + ox) * p.out_pitch + co;
    dst[p.out_hi_off] = hi;
    if (p.write_lo) dst[p.out_lo_off] = lo;
  }
}

}  // namespace b200al
