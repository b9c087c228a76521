// Implicit-GEMM convolution on tcgen05 tensor cores (sm_100a).
//
// One kernel covers the UNet's 3x3 convs (stride 1 and 2), and the 2x2/stride-2
// transposed convs (as a 1-tap GEMM with a pixel-shuffle store):
//   D[128 pixels, BLOCK_N couts] = sum over (tap, cin-chunk) A_tap[128, BK] * W_tap[BLOCK_N, BK]^T
// * A tiles are im2col-free: a 5-D TMA box (BK channels x tw x 1 x th x 1) of the
//   NHWC activation tensor per filter tap; conv zero padding = TMA out-of-bounds fill.
//   Stride-2 convs view the input as (2*pitch, W/2, 2, H/2, S) so each tap is a
//   plain box (no element strides).
// * Operands are bf16 "hi/lo" pairs (x ~= hi + lo, 16 significand bits); with
//   NSPLIT==3 every k-step issues hi*hi + hi*lo + lo*hi into one fp32 TMEM
//   accumulator (parity-grade, the default); NSPLIT==1 uses the hi planes only.
// * Persistent CTAs: grid = min(tiles, SMs x CTAs/SM); every CTA walks tiles
//   blockIdx.x, +gridDim.x, ...  Warp roles: warp 0 = TMA producer, warp 1 =
//   UMMA issuer, warps 2..5 = epilogue (TMEM -> registers -> folded BN affine
//   (+MC-dropout mask) -> LeakyReLU -> hi/lo split -> NHWC stores).  The smem
//   ring (full/empty mbarriers) runs across tile boundaries and the fp32
//   accumulator is double-buffered in TMEM (tmem_full/tmem_empty mbarriers),
//   so the loads and MMAs of tile i+1 overlap the epilogue of tile i.
#pragma once

#include "common.cuh"

namespace b200al {

struct ConvTcParams {
  // ---- k loop -------------------------------------------------------------
  int taps;        // 9 (3x3 conv) or 1 (plain GEMM / transposed conv)
  int chunks;      // Cin / BK
  int a_c0[3];     // inner-dim (channel) base coordinate per tap column s
  int a_dx[3];     // x offset per tap column s
  int a_p[3];      // row-parity coordinate per tap row r (stride-2 view), else 0
  int a_dy[3];     // y offset per tap row r
  int a_lo_off;    // inner-dim offset of the lo plane
  int n_total;     // GEMM N (Cout, or 4*Cout for the transposed conv)
  // ---- tiling over the image the A tiles walk (conv: output grid) ----------
  int tw, th, tiles_x, tiles_y;
  int n_tiles;      // n_total / BLOCK_N
  int total_tiles;  // samples * tiles_y * tiles_x * n_tiles
  // ---- epilogue -------------------------------------------------------------
  const float* scale;  // [cout]  gamma / sqrt(var + eps)        (1 for convT)
  const float* bias;   // [cout]  conv bias
  const float* shift;  // [cout]  beta - mean * scale            (0 for convT)
  const float* mask;   // nullable: per-sample multipliers, mask[s*mask_stride + c]
  int mask_stride;
  float slope;         // LeakyReLU slope (1 = identity)
  __nv_bfloat16* out;
  long long out_pitch;  // elements per output pixel
  int out_hi_off, out_lo_off;
  int out_W, out_H;
  int shuffle;  // 1: transposed conv, column n -> (quad=n/cout, co=n%cout), pixel (2y+quad/2, 2x+quad%2)
  int cout;     // channels per quad (== n_total when !shuffle)
  int write_lo;
};

constexpr int kConvThreads = 192;

template <int BLOCK_N, int BK, int NSPLIT>
struct ConvTcSmem {
  static constexpr int kPlanes = (NSPLIT == 3) ? 2 : 1;
  static constexpr int kABytes = 128 * BK * 2;
  static constexpr int kBBytes = BLOCK_N * BK * 2;
  static constexpr int kStageBytes = kPlanes * (kABytes + kBBytes);
  static constexpr int kTailBytes = 1024;  // barriers, tmem ptr, epilogue affine
  static constexpr int bytes(int stages) { return stages * kStageBytes + kTailBytes + 4 * BLOCK_N * 4 + 1024; }
};

struct TileCoord {
  int sample, x0, y0, n0;
};
__device__ __forceinline__ TileCoord decode_tile(const ConvTcParams& p, int t, int block_n) {
  TileCoord c;
  const int nt = t % p.n_tiles;
  int m = t / p.n_tiles;
  const int tiles_per_sample = p.tiles_x * p.tiles_y;
  c.sample = m / tiles_per_sample;
  m -= c.sample * tiles_per_sample;
  const int ty = m / p.tiles_x;
  c.x0 = (m - ty * p.tiles_x) * p.tw;
  c.y0 = ty * p.th;
  c.n0 = nt * block_n;
  return c;
}

template <int BLOCK_N, int BK, int NSPLIT>
__global__ void __launch_bounds__(kConvThreads)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const ConvTcParams p, const int stages) {
  using L = ConvTcSmem<BLOCK_N, BK, NSPLIT>;
  constexpr uint32_t kRowBytes = BK * 2;  // 128 -> SW128, 64 -> SW64
  constexpr uint32_t kTmemCols = 2 * BLOCK_N;  // two accumulator stages (power of two >= 64)
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled tiles
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tail = smem + (size_t)stages * L::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + 16;
  uint64_t* tfull_bar = empty_bar + 16;   // [2] accumulator ready
  uint64_t* tempty_bar = tfull_bar + 2;   // [2] accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* es = reinterpret_cast<float*>(tail + L::kTailBytes);  // [2][BLOCK_N]
  float* eh = es + 2 * BLOCK_N;                                // [2][BLOCK_N]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = p.taps * p.chunks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tfull_bar[i]), 1);
      mbar_init(smem_u32(&tempty_bar[i]), 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(p, t, BLOCK_N);
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / p.chunks;
          const int ch = kb - tap * p.chunks;
          const int r = tap / 3, s = tap - r * 3;
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_arrive_expect_tx(fb, L::kStageBytes);
          uint8_t* st = smem + (size_t)stage * L::kStageBytes;
          const int c = p.a_c0[s] + ch * BK;
          const int cx = tc.x0 + p.a_dx[s], cy = tc.y0 + p.a_dy[r], cp = p.a_p[r];
          tma_load_5d(smem_u32(st), &tmA, fb, c, cx, cp, cy, tc.sample);
          tma_load_2d(smem_u32(st + L::kPlanes * L::kABytes), &tmB, fb, ch * BK, tap * p.n_total + tc.n0);
          if (NSPLIT == 3) {
            tma_load_5d(smem_u32(st + L::kABytes), &tmA, fb, c + p.a_lo_off, cx, cp, cy, tc.sample);
            tma_load_2d(smem_u32(st + 2 * L::kABytes + L::kBBytes), &tmB, fb, ch * BK,
                        (p.taps + tap) * p.n_total + tc.n0);
          }
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t a_hi = smem_u32(smem + (size_t)stage * L::kStageBytes);
          const uint32_t a_lo = a_hi + L::kABytes;
          const uint32_t b_hi = a_hi + L::kPlanes * L::kABytes;
          const uint32_t b_lo = b_hi + L::kBBytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = umma_smem_desc(a_hi + k * 32, kRowBytes);
            const uint64_t db = umma_smem_desc(b_hi + k * 32, kRowBytes);
            umma_bf16(d_tmem, da, db, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            if (NSPLIT == 3) {
              umma_bf16(d_tmem, da, umma_smem_desc(b_lo + k * 32, kRowBytes), idesc, 1u);
              umma_bf16(d_tmem, umma_smem_desc(a_lo + k * 32, kRowBytes), db, idesc, 1u);
            }
          }
          umma_commit(smem_u32(&empty_bar[stage]));  // frees the smem slot when the MMAs retire
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));  // accumulator complete
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int et = threadIdx.x - 64;  // 0..127
    const int q = warp & 3;           // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;
    const int ry = row / p.tw, rx = row - ry * p.tw;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
      const TileCoord tc = decode_tile(p, t, BLOCK_N);
      float* esa = es + acc * BLOCK_N;
      float* eha = eh + acc * BLOCK_N;
      for (int i = et; i < BLOCK_N; i += 128) {
        const int n = tc.n0 + i;
        const int c = p.shuffle ? (n % p.cout) : n;
        float sc = p.scale[c];
        if (p.mask) sc *= p.mask[(size_t)tc.sample * p.mask_stride + c];
        esa[i] = sc;
        eha[i] = p.bias[c] * sc + p.shift[c];
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BLOCK_N + c0), v);
        tmem_ld_wait();
        if (c0 + 32 >= BLOCK_N) {  // last read of this accumulator: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
        }
        int oy = tc.y0 + ry, ox = tc.x0 + rx, cbase = tc.n0 + c0;
        if (p.shuffle) {
          const int quad = cbase / p.cout;
          cbase -= quad * p.cout;
          oy = 2 * oy + (quad >> 1);
          ox = 2 * ox + (quad & 1);
        }
        __nv_bfloat16* dst =
            p.out + ((size_t)((size_t)tc.sample * p.out_H + oy) * p.out_W + ox) * p.out_pitch + cbase;
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float y0f = leaky(__uint_as_float(v[2 * j]) * esa[c0 + 2 * j] + eha[c0 + 2 * j], p.slope);
          const float y1f =
              leaky(__uint_as_float(v[2 * j + 1]) * esa[c0 + 2 * j + 1] + eha[c0 + 2 * j + 1], p.slope);
          __nv_bfloat16 h0, l0, h1, l1;
          split_bf16(y0f, h0, l0);
          split_bf16(y1f, h1, l1);
          hi[j] = pack_bf16x2(h0, h1);
          lo[j] = pack_bf16x2(l0, l1);
        }
        uint4* dh = reinterpret_cast<uint4*>(dst + p.out_hi_off);
#pragma unroll
        for (int j = 0; j < 4; ++j) dh[j] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
        if (p.write_lo) {
          uint4* dl = reinterpret_cast<uint4*>(dst + p.out_lo_off);
#pragma unroll
          for (int j = 0; j < 4; ++j) dl[j] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace b200al
