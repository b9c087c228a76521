// Column-folded 3x3 convolution on tcgen05 for the wide, shallow layers
// (stride 1, Cout = 32, Cin in {32, 64}: the full-resolution UNet blocks).
//
// The generic kernel (conv_tc.cuh) fetches one A tile per filter tap, i.e. it
// pulls every activation 9x from L2 and issues N = 32 MMAs whose shared-memory
// operand reads outweigh their math.  Here the three column taps s are folded
// into the GEMM N dimension and the three row taps r are resolved inside one
// shared-memory halo tile:
//
//   D[pixel (y,x), (s, co)] = sum_{r, ci} in[y+r-1, x, ci] * W[r][s][co][ci]          (N = 96, K = 3*Cin)
//   out[y, x, co]           = D[(y,x-1), (0,co)] + D[(y,x), (1,co)] + D[(y,x+1), (2,co)]
//
// * One TMA box of 6 rows x 32 columns x Cin per plane feeds all three row taps:
//   tap r is the same box entered 32*r rows further down (a whole number of
//   8-row swizzle groups), so the UMMA descriptor only moves its start address.
// * A tile is 4 image rows x 32 columns (one warp quarter = one image row); the
//   x-1 / x+1 terms are warp shuffles in the epilogue, so columns 0 and 31 of a
//   tile are halo: a tile yields 30 output columns and tiles step by 30.
// * The packed weights of the layer (both planes, 36-72 KB) stay resident in
//   shared memory for the life of the persistent CTA.
// * L2->SM traffic per output pixel drops from 9 to 1.7 activation reads, the
//   MMA count by 3x at 86% instead of 40% issue efficiency.
// Operand precision, epilogue math and the output format are those of conv_tc.cuh.
#pragma once

#include "common.cuh"

namespace b200al {

struct ConvFoldParams {
  int a_c0;       // channel offset of the hi plane inside the input pixel
  int a_lo_off;   // + offset of the lo plane
  int tiles_x;    // ceil(W / 30)
  int tiles_y;    // H / 4
  int total_tiles;
  const float* scale;
  const float* bias;
  const float* shift;
  const float* mask;  // nullable, mask[s * mask_stride + c]
  int mask_stride;
  float slope;
  int out_hi_off, out_lo_off;
  // fused 1x1 seg_output head (HEADK > 0): logits [S][HEADK][H][W] fp32 instead of the activation
  const float* head_w;  // [HEADK][32]
  const float* head_b;  // [HEADK]
  float* logits;
  int H, W;
  int w_cout_total;  // output channels of the whole layer (row stride between filter taps in the packed weights)
  int w_co0;         // first output channel of the 32 this launch computes
  int dbg;  // timing experiments only (B200AL_FOLD_DBG): 1 = no MMAs, 2 = no epilogue math/stores, 4 = no TMA stores, 8 = no TMA loads, 16 = print the MMA warp's loop cycles
};

constexpr int kFoldThreads = 320;   // TMA warp, MMA warp, 2 epilogue groups of 4 warps
constexpr int kFoldN = 96;          // 3 column taps x 32 output channels
constexpr int kFoldValid = 30;      // output columns per tile
// TMEM columns between the two accumulators.  Stacked (CIN == 64, bf16x3): a_hi * [w_hi ; w_lo] is ONE MMA of
// N = 192 into a double-width accumulator (the hi and lo weight blocks of a filter row are adjacent in shared
// memory; 96 cycles, all math, instead of 2 x 56 shared-memory-read-bound cycles), a_lo * w_hi adds into its first
// 96 columns and the epilogue sums the halves.
__host__ __device__ constexpr int fold_acc_stride(bool stack) { return stack ? 256 : 128; }

// Each persistent CTA owns a contiguous range of tiles (sample-major, then tile row, then tile
// column) and steps through it without divisions.
struct FoldTileWalk {
  int t, t_end, sample, ty, tx;
  __device__ __forceinline__ void init(const ConvFoldParams& p) {
    const long long total = p.total_tiles;
    t = (int)(total * blockIdx.x / gridDim.x);
    t_end = (int)(total * (blockIdx.x + 1) / gridDim.x);
    const int tps = p.tiles_x * p.tiles_y;
    sample = t / tps;
    const int rem = t - sample * tps;
    ty = rem / p.tiles_x;
    tx = rem - ty * p.tiles_x;
  }
  __device__ __forceinline__ bool valid() const { return t < t_end; }
  __device__ __forceinline__ void next(const ConvFoldParams& p) {
    ++t;
    if (++tx == p.tiles_x) {
      tx = 0;
      if (++ty == p.tiles_y) { ty = 0; ++sample; }
    }
  }
};

// CIN == 32 ("packed"): the hi and lo planes of a 32-channel pixel are adjacent in memory, so one
// 128-byte SWIZZLE_128B row carries both (hi in bytes 0-63, lo in 64-127) and the weights are packed
// the same way; the tensor core fetches 128-byte rows noticeably faster than 64-byte ones.
// CIN == 64: one 128-byte row per plane, the planes travel in separate ring slots.
template <int CIN, int NSPLIT>
struct ConvFoldSmem {
  static constexpr bool kPacked = (CIN == 32);
  static constexpr int kPlanes = (NSPLIT == 3) ? 2 : 1;
  static constexpr int kRowBytes = kPacked ? 128 : CIN * 2;
  static constexpr int kWBlockBytes = kFoldN * kRowBytes;        // one filter row r (packed: both planes)
  static constexpr int kWBytes = (kPacked ? 1 : kPlanes) * 3 * kWBlockBytes;
  static constexpr int kSlotBytes = 6 * 32 * kRowBytes;          // one halo tile (packed) / one plane of it
  static constexpr int kStagePlaneBytes = 2048;                  // 30 px x 64 B, padded
  static constexpr int kStagingBytes = 8 * kPlanes * kStagePlaneBytes;
  static constexpr int kTailBytes = 3072;                        // barriers, tmem ptr, affine + head tables
  static constexpr int bytes(int slots) { return kWBytes + slots * kSlotBytes + kStagingBytes + kTailBytes; }
};

// HEADK > 0 (last decoder block only): the epilogue applies the UNet's 1x1 seg_output conv to the
// 32 fp32 channel values it holds in registers and stores HEADK logits per pixel; the activation
// itself is not written.
template <int CIN, int NSPLIT, int HEADK>
__global__ void __launch_bounds__(kFoldThreads)
conv_fold_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmO, const ConvFoldParams p, const int slots) {
  using L = ConvFoldSmem<CIN, NSPLIT>;
  constexpr uint32_t kRowBytes = L::kRowBytes;
  constexpr bool STACK = !L::kPacked && NSPLIT == 3;
  constexpr uint32_t kFoldAccStride = fold_acc_stride(STACK);
  constexpr uint32_t kTmemCols = 2 * kFoldAccStride;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* wsm = smem_raw;
  uint8_t* ring = wsm + L::kWBytes;
  uint8_t* staging = ring + (size_t)slots * L::kSlotBytes;
  uint8_t* tail = staging + L::kStagingBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);   // [16]
  uint64_t* empty_bar = full_bar + 16;                       // [16]
  uint64_t* tfull_bar = empty_bar + 16;                      // [2]
  uint64_t* tempty_bar = tfull_bar + 2;                      // [2]
  uint64_t* w_bar = tempty_bar + 2;                          // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);
  float* es = reinterpret_cast<float*>(tail + 512);          // [2 groups][32]
  float* eh = es + 64;                                       // [2 groups][32]
  float* hws = reinterpret_cast<float*>(tail + 1024);        // [HEADK][32] head weights, then [HEADK] bias

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem_raw) & 1023u) {
      printf("b200al: dynamic shared memory is not 1024-byte aligned\n");
      __trap();
    }
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < slots; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tfull_bar[i]), 1);
      mbar_init(smem_u32(&tempty_bar[i]), 4);
    }
    mbar_init(smem_u32(w_bar), 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  if (HEADK > 0 && warp >= 3) {
    for (int i = threadIdx.x - 96; i < HEADK * 33; i += kFoldThreads - 96)
      hws[i] = i < HEADK * 32 ? p.head_w[i] : p.head_b[i - HEADK * 32];
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      // resident weights: blocks of 96 rows (s, co), one per filter row r (and per plane unless packed)
      const uint32_t wb = smem_u32(w_bar);
      mbar_arrive_expect_tx(wb, L::kWBytes);
      if (L::kPacked) {
        for (int r = 0; r < 3; ++r) tma_load_2d(smem_u32(wsm + r * L::kWBlockBytes), &tmW, wb, 0, r * 3 * 32);
      } else {
        for (int pl = 0; pl < L::kPlanes; ++pl)
          for (int r = 0; r < 3; ++r)
            for (int s3 = 0; s3 < 3; ++s3)  // 32 of the layer's output channels per column tap
              tma_load_2d(smem_u32(wsm + (r * L::kPlanes + pl) * L::kWBlockBytes + s3 * 32 * kRowBytes), &tmW, wb, 0,
                          (pl * 9 + r * 3 + s3) * p.w_cout_total + p.w_co0);
      }
      int slot = 0;
      uint32_t phase = 0;
      FoldTileWalk w;
      for (w.init(p); w.valid(); w.next(p)) {
        const int cx = w.tx * kFoldValid - 1, cy = w.ty * 4 - 1;
        for (int pl = 0; pl < (L::kPacked ? 1 : L::kPlanes); ++pl) {
          mbar_wait(smem_u32(&empty_bar[slot]), phase ^ 1);
          const uint32_t fb = smem_u32(&full_bar[slot]);
          if (p.dbg & 8) {  // timing experiment: no activation loads at all
            mbar_arrive(fb);
          } else {
            mbar_arrive_expect_tx(fb, L::kSlotBytes);
            tma_load_5d(smem_u32(ring + (size_t)slot * L::kSlotBytes), &tmA, fb, p.a_c0 + pl * p.a_lo_off, cx, 0, cy,
                        w.sample);
          }
          if (++slot == slots) { slot = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer =====================
    // The tensor pipe runs dry whenever the issuing warp spends more than about one MMA's time away
    // from issuing (tools/mma_bench.cu), so the bookkeeping is spread between the MMAs: the whole warp
    // walks the loop (descriptor arithmetic stays on the uniform datapath), an elected lane issues,
    // and the NEXT tile's barriers are polled and its descriptors built half-way through this tile.
    constexpr uint32_t idesc = umma_idesc_bf16(128, kFoldN);
    constexpr uint32_t idesc2 = umma_idesc_bf16(128, 2 * kFoldN);
    constexpr uint32_t kDescHi = umma_smem_desc_hi(kRowBytes);
    constexpr int KS = L::kPacked ? 2 : CIN / 16;   // k16 steps per filter row and plane
    constexpr int NA = 3 * KS;                      // A descriptors per plane
    mbar_wait(smem_u32(w_bar), 0);
    tc_fence_after();
    const uint32_t w_base = umma_smem_desc_lo(smem_u32(wsm));
    const uint32_t ring_lo = umma_smem_desc_lo(smem_u32(ring));
    // weight descriptors never change
    uint64_t dwh[NA], dwl[NA];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int k = 0; k < KS; ++k) {
        // packed: [r][96 rows of hi 64 B | lo 64 B]; otherwise [r][plane][96 rows]
        const uint32_t wo = w_base + (((L::kPacked ? r : r * L::kPlanes) * L::kWBlockBytes + k * 32) >> 4);
        dwh[r * KS + k] = make_desc(wo, kDescHi);
        dwl[r * KS + k] = make_desc(wo + (L::kPacked ? 4u : (uint32_t)(L::kWBlockBytes >> 4)), kDescHi);
      }
    const int n_my = (int)((long long)p.total_tiles * (blockIdx.x + 1) / gridDim.x) -
                     (int)((long long)p.total_tiles * blockIdx.x / gridDim.x);
    const bool mma_on = !(p.dbg & 1);
    const long long t_start = clock64();
    int slot = 0;
    uint32_t phase = 0;
    uint32_t acc = 0, acc_phase = 0;
    bool ready = false;
    auto a_descs = [&](uint64_t (&d)[NA], int sl, uint32_t extra) {
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int k = 0; k < KS; ++k)
          d[r * KS + k] = make_desc(ring_lo + (uint32_t)sl * (L::kSlotBytes >> 4) + extra +
                                        ((r * 32 * kRowBytes + k * 32) >> 4), kDescHi);
    };
    if (L::kPacked) {
      uint64_t dah[NA], dal[NA];
      a_descs(dah, 0, 0);
      a_descs(dal, 0, 4);  // lo plane: +64 bytes inside the row
      for (int i = 0; i < n_my; ++i) {
        if (!ready) {
          mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1);
          mbar_wait(smem_u32(&full_bar[slot]), phase);
        }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kFoldAccStride;
        if (mma_on && elect_one()) {
#pragma unroll
          for (int m = 0; m < NA / 2; ++m) {
            umma_bf16(d_tmem, dah[m], dwh[m], idesc, m > 0 ? 1u : 0u);
            if (NSPLIT == 3) {
              umma_bf16(d_tmem, dah[m], dwl[m], idesc, 1u);
              umma_bf16(d_tmem, dal[m], dwh[m], idesc, 1u);
            }
          }
        }
        int nslot = slot + 1;
        uint32_t nphase = phase;
        if (nslot == slots) { nslot = 0; nphase ^= 1; }
        const uint32_t nacc = acc ^ 1, nacc_phase = acc_phase ^ (nacc == 0 ? 1u : 0u);
        ready = false;
        if (i + 1 < n_my)
          ready = mbar_test_wait(smem_u32(&tempty_bar[nacc]), nacc_phase ^ 1) &&
                  mbar_test_wait(smem_u32(&full_bar[nslot]), nphase);
        uint64_t nah[NA], nal[NA];
        a_descs(nah, nslot, 0);
        a_descs(nal, nslot, 4);
        if (elect_one()) {
          if (mma_on) {
#pragma unroll
            for (int m = NA / 2; m < NA; ++m) {
              umma_bf16(d_tmem, dah[m], dwh[m], idesc, 1u);
              if (NSPLIT == 3) {
                umma_bf16(d_tmem, dah[m], dwl[m], idesc, 1u);
                umma_bf16(d_tmem, dal[m], dwh[m], idesc, 1u);
              }
            }
          }
          umma_commit(smem_u32(&empty_bar[slot]));
          umma_commit(smem_u32(&tfull_bar[acc]));
        }
#pragma unroll
        for (int m = 0; m < NA; ++m) { dah[m] = nah[m]; dal[m] = nal[m]; }
        slot = nslot; phase = nphase; acc = nacc; acc_phase = nacc_phase;
      }
    } else {
      // planes in separate slots: hi slot first (a_hi*w_hi, a_hi*w_lo), then lo slot (a_lo*w_hi)
      uint64_t dah[NA];
      a_descs(dah, 0, 0);
      for (int i = 0; i < n_my; ++i) {
        if (!ready) {
          mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1);
          mbar_wait(smem_u32(&full_bar[slot]), phase);
        }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kFoldAccStride;
        if (mma_on && elect_one()) {
#pragma unroll
          for (int m = 0; m < NA / 2; ++m) {
            if (STACK) {
              umma_bf16(d_tmem, dah[m], dwh[m], idesc2, m > 0 ? 1u : 0u);  // [hi*hi | hi*lo]
            } else {
              umma_bf16(d_tmem, dah[m], dwh[m], idesc, m > 0 ? 1u : 0u);
              if (NSPLIT == 3) umma_bf16(d_tmem, dah[m], dwl[m], idesc, 1u);
            }
          }
        }
        // second slot of this tile (lo plane) or, in bf16 mode, the next tile's slot
        int slot1 = slot + 1;
        uint32_t phase1 = phase;
        if (slot1 == slots) { slot1 = 0; phase1 ^= 1; }
        const uint32_t nacc = acc ^ 1, nacc_phase = acc_phase ^ (nacc == 0 ? 1u : 0u);
        bool ready1 = false;
        if (NSPLIT == 3) ready1 = mbar_test_wait(smem_u32(&full_bar[slot1]), phase1);
        else if (i + 1 < n_my)
          ready1 = mbar_test_wait(smem_u32(&tempty_bar[nacc]), nacc_phase ^ 1) &&
                   mbar_test_wait(smem_u32(&full_bar[slot1]), phase1);
        uint64_t d1[NA];
        a_descs(d1, slot1, 0);
        if (elect_one()) {
          if (mma_on) {
#pragma unroll
            for (int m = NA / 2; m < NA; ++m) {
              if (STACK) {
                umma_bf16(d_tmem, dah[m], dwh[m], idesc2, 1u);
              } else {
                umma_bf16(d_tmem, dah[m], dwh[m], idesc, 1u);
                if (NSPLIT == 3) umma_bf16(d_tmem, dah[m], dwl[m], idesc, 1u);
              }
            }
          }
          umma_commit(smem_u32(&empty_bar[slot]));
          if (NSPLIT == 1) umma_commit(smem_u32(&tfull_bar[acc]));
        }
        if (NSPLIT == 1) {
#pragma unroll
          for (int m = 0; m < NA; ++m) dah[m] = d1[m];
          ready = ready1;
          slot = slot1; phase = phase1; acc = nacc; acc_phase = nacc_phase;
          continue;
        }
        if (!ready1) mbar_wait(smem_u32(&full_bar[slot1]), phase1);
        tc_fence_after();
        if (mma_on && elect_one()) {
#pragma unroll
          for (int m = 0; m < NA / 2; ++m) umma_bf16(d_tmem, d1[m], dwh[m], idesc, 1u);
        }
        int nslot = slot1 + 1;
        uint32_t nphase = phase1;
        if (nslot == slots) { nslot = 0; nphase ^= 1; }
        ready = false;
        if (i + 1 < n_my)
          ready = mbar_test_wait(smem_u32(&tempty_bar[nacc]), nacc_phase ^ 1) &&
                  mbar_test_wait(smem_u32(&full_bar[nslot]), nphase);
        a_descs(dah, nslot, 0);
        if (elect_one()) {
          if (mma_on) {
#pragma unroll
            for (int m = NA / 2; m < NA; ++m) umma_bf16(d_tmem, d1[m], dwh[m], idesc, 1u);
          }
          umma_commit(smem_u32(&empty_bar[slot1]));
          umma_commit(smem_u32(&tfull_bar[acc]));
        }
        slot = nslot; phase = nphase; acc = nacc; acc_phase = nacc_phase;
      }
    }
    if ((p.dbg & 16) && blockIdx.x == 0 && lane == 0)
      printf("b200al fold dbg: %d tiles, %lld cycles in the issue loop\n", n_my, clock64() - t_start);
  } else {
    // ===================== epilogue: two groups of 4 warps, alternating tiles =====================
    const int group = (warp - 2) >> 2;
    const int et = (threadIdx.x - 64) & 127;
    const int q = warp & 3;  // TMEM lane quarter == image row of the tile
    uint8_t* st_hi = staging + (size_t)(group * 4 + q) * L::kPlanes * L::kStagePlaneBytes;
    uint8_t* st_lo = st_hi + L::kStagePlaneBytes;
    float* esa = es + group * 32;
    float* eha = eh + group * 32;
    int prev_sample = -1;
    int it = 0;
    FoldTileWalk w;
    for (w.init(p); w.valid(); w.next(p), ++it) {
      if ((it & 1) != group) continue;
      const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
      const int sample = w.sample, ty = w.ty, tx = w.tx;
      if (sample != prev_sample) {  // uniform across the group
        prev_sample = sample;
        if (group == 0) asm volatile("bar.sync 1, 128;" ::: "memory");  // previous tile's table reads are done
        else asm volatile("bar.sync 2, 128;" ::: "memory");
        if (et < 32) {
          float sc = p.scale[et];
          if (p.mask) sc *= p.mask[(size_t)sample * p.mask_stride + et];
          esa[et] = sc;
          eha[et] = p.bias[et] * sc + p.shift[et];
        }
        if (group == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
        else asm volatile("bar.sync 2, 128;" ::: "memory");
      }
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
      if (p.dbg & 2) {
        tc_fence_before();
        __syncwarp();
        if (elect_one()) mbar_arrive(smem_u32(&tempty_bar[acc]));
        continue;
      }
      // the staging rows must no longer be read by this warp's previous TMA store
      if (HEADK == 0) {
        if (elect_one()) tma_store_wait_read<0>();
        __syncwarp();
      }
      float lg[HEADK > 0 ? HEADK : 1];
#pragma unroll
      for (int k = 0; k < HEADK; ++k) lg[k] = hws[HEADK * 32 + k];
      const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + acc * kFoldAccStride;
      const int row = lane - 1;  // staging row of this lane's output pixel (lanes 1..30)
      // pull the whole accumulator into registers and hand the TMEM stage back before the math
      uint32_t dd[2][3][16];
      if (!STACK) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
          for (int s3 = 0; s3 < 3; ++s3) tmem_ld_32x32_x16(trow + s3 * 32 + h * 16, dd[h][s3]);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (elect_one()) mbar_arrive(smem_u32(&tempty_bar[acc]));
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {  // 16 output channels per pass
        if (STACK) {  // both halves of the stacked accumulator, 16 channels at a time (register budget)
          uint32_t e2[3][16];
#pragma unroll
          for (int s3 = 0; s3 < 3; ++s3) {
            tmem_ld_32x32_x16(trow + s3 * 32 + h * 16, dd[h][s3]);
            tmem_ld_32x32_x16(trow + kFoldN + s3 * 32 + h * 16, e2[s3]);
          }
          tmem_ld_wait();
          if (h == 1) {
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive(smem_u32(&tempty_bar[acc]));
          }
#pragma unroll
          for (int s3 = 0; s3 < 3; ++s3)
#pragma unroll
            for (int j = 0; j < 16; ++j)
              dd[h][s3][j] = __float_as_uint(__uint_as_float(dd[h][s3][j]) + __uint_as_float(e2[s3][j]));
        }
        const uint32_t(&d0)[16] = dd[h][0];
        const uint32_t(&d1)[16] = dd[h][1];
        const uint32_t(&d2)[16] = dd[h][2];
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int cc = h * 16 + 2 * j;
          const float l0 = __shfl_up_sync(0xffffffffu, __uint_as_float(d0[2 * j]), 1);
          const float l1 = __shfl_up_sync(0xffffffffu, __uint_as_float(d0[2 * j + 1]), 1);
          const float r0 = __shfl_down_sync(0xffffffffu, __uint_as_float(d2[2 * j]), 1);
          const float r1 = __shfl_down_sync(0xffffffffu, __uint_as_float(d2[2 * j + 1]), 1);
          const float a0 = (l0 + __uint_as_float(d1[2 * j])) + r0;
          const float a1 = (l1 + __uint_as_float(d1[2 * j + 1])) + r1;
          const float2 sc2 = *reinterpret_cast<const float2*>(esa + cc);
          const float2 sh2 = *reinterpret_cast<const float2*>(eha + cc);
          float y0f = fmaf(a0, sc2.x, sh2.x);
          float y1f = fmaf(a1, sc2.y, sh2.y);
          y0f = fmaxf(y0f, y0f * p.slope);
          y1f = fmaxf(y1f, y1f * p.slope);
          if (HEADK > 0) {
#pragma unroll
            for (int k = 0; k < HEADK; ++k) {
              const float2 w2 = *reinterpret_cast<const float2*>(hws + k * 32 + cc);
              lg[k] = fmaf(y1f, w2.y, fmaf(y0f, w2.x, lg[k]));
            }
            continue;
          }
          const __nv_bfloat162 h2 = __floats2bfloat162_rn(y0f, y1f);
          const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h2);
          hi[j] = hb;
          if (NSPLIT == 3) {
            const __nv_bfloat162 l2 =
                __floats2bfloat162_rn(y0f - __uint_as_float(hb << 16), y1f - __uint_as_float(hb & 0xffff0000u));
            lo[j] = *reinterpret_cast<const uint32_t*>(&l2);
          }
        }
        if (HEADK == 0 && lane >= 1 && lane <= kFoldValid) {
          // 64-byte rows, SWIZZLE_64B: 16-byte chunk c of row `row` lives at chunk (c ^ ((row >> 1) & 3))
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int chunk = h * 2 + j;
            const uint32_t off = (uint32_t)(row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4));
            *reinterpret_cast<uint4*>(st_hi + off) = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
            if (NSPLIT == 3)
              *reinterpret_cast<uint4*>(st_lo + off) =
                  make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
          }
        }
      }
      const int ox = tx * kFoldValid, oy = ty * 4 + q;
      if (HEADK > 0) {
        const int x = ox + lane - 1;
        if (lane >= 1 && lane <= kFoldValid && x < p.W) {
          float* dst = p.logits + (((size_t)sample * HEADK) * p.H + oy) * p.W + x;
#pragma unroll
          for (int k = 0; k < HEADK; ++k) dst[(size_t)k * p.H * p.W] = lg[k];
        }
        continue;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (!(p.dbg & 4) && elect_one()) {
        tma_store_5d(&tmO, smem_u32(st_hi), p.out_hi_off, ox, 0, oy, sample);
        if (NSPLIT == 3) tma_store_5d(&tmO, smem_u32(st_lo), p.out_lo_off, ox, 0, oy, sample);
        tma_store_commit();
      }
    }
    if (HEADK == 0 && elect_one()) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace b200al
