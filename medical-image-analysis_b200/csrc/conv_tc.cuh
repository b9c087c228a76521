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
  // 1: launched as clusters of 2 CTAs (template PAIR) that run one M=256 cta_group::2 MMA over two adjacent
  // pixel tiles of the same n-tile; each CTA loads its own pixel tile and HALF of the weight tile
  int cluster;
  // 2: every CTA walks SUPER-tiles of two adjacent pixel tiles that share one weight tile in shared memory
  // (template MT == 2): the weight bytes per MMA halve without any cross-CTA traffic.  1 otherwise.
  int mt;
};

constexpr int kConvThreads = 320;  // TMA warp + MMA warp + 2 epilogue groups of 4 warps

// FMT (mixed-precision scoring path): 0 = operands and output as NSPLIT says (bf16 hi/lo pair, or hi plane only);
// 1 = bf16 hi/lo operands (NSPLIT == 3), output ONE fp16 plane (the layer in front of an fp16 group);
// 2 = single fp16 operands (NSPLIT == 1, one MMA per k-step, kind::f16 with f16 formats), output one fp16 plane;
// 3 = single fp16 operands, output a bf16 hi/lo pair again (last layer of an fp16 group).
// A single plane always sits where the hi plane of the pair would (the lo half of the pixel stays unused).
// smem: [stages x (A planes, B planes)] [nbuf x planes x 128 x CW staging for the TMA store] [barriers] [affine]
// HALO (fp16 stride-1 3x3 layers): a pipeline stage holds ONE activation box of 10 image rows x 16 pixels x BK
// channels for a filter column s, and the weight tiles of the three filter rows r: tap (r, s) reads the box from image
// row r on (whole 1024-byte swizzle groups further down, only the descriptor start address moves).  A bytes per MMA
// drop from 9 boxes of 8 rows to 3 boxes of 10, a stage carries 3 x BK/16 MMAs per barrier hand-shake.
template <int BLOCK_N, int BK, int NSPLIT, int CW, bool PAIR = false, int MT = 1, int FMT = 0, bool HALO = false>
struct ConvTcSmem {
  // operand planes per pipeline stage: the hi/lo pair (NSPLIT == 3), or - fp16 layers (FMT >= 2) - TWO consecutive
  // BK-channel chunks, so that one barrier hand-shake feeds 2 x BK/16 MMAs instead of BK/16 (with a single MMA per
  // k-step the hand-shake latency, not the pipe, bounded these layers: profiles/r02_ncu_full_conv.csv)
  static constexpr int kKChunks = (FMT >= 2 && !HALO) ? 2 : 1;
  static constexpr int kPlanes = (NSPLIT == 3) ? 2 : kKChunks;
  static constexpr int kOutPlanes = FMT == 0 ? ((NSPLIT == 3) ? 2 : 1) : (FMT == 3 ? 2 : 1);
  static constexpr int kABytes = (HALO ? 160 : 128) * BK * 2;
  static constexpr int kBBytes = (PAIR ? BLOCK_N / 2 : BLOCK_N) * BK * 2;  // rows of B held by this CTA
  static constexpr int kBTaps = HALO ? 3 : 1;                              // weight tiles per stage
  // [A tile(s): hi, lo per tile][B: hi, lo]   (HALO: [A box][B row 0][B row 1][B row 2])
  static constexpr int kStageBytes = kPlanes * (MT * kABytes + kBTaps * kBBytes);
  static constexpr int kOutWarpBytes = 32 * CW * 2;                  // one warp's 32 rows of one plane
  static constexpr int kOutBufBytes = kOutPlanes * 4 * kOutWarpBytes;   // one staging buffer
  static constexpr int kTailBytes = 1024;  // barriers, tmem ptr
};

struct TileCoord {
  int sample, x0, y0, n0;
};
// i-th tile of this CTA.  Plain launch: tiles blockIdx.x, +gridDim.x, ... with the n-tile fastest.
// Cluster launch: cluster c walks tile PAIRS c, +nclusters, ...; a pair is two consecutive pixel tiles
// of one n-tile, CTA rank r takes the r-th.
__device__ __forceinline__ bool get_tile(const ConvTcParams& p, int i, int block_n, uint32_t rank, TileCoord* c) {
  int nt, m;
  if (p.cluster) {
    const int pair = (int)(blockIdx.x >> 1) + i * (int)(gridDim.x >> 1);
    if (pair >= (p.total_tiles >> 1)) return false;
    nt = pair % p.n_tiles;
    m = (pair / p.n_tiles) * 2 + (int)rank;
  } else if (p.mt == 2) {  // rank = which of the super-tile's two pixel tiles
    const int pair = (int)blockIdx.x + i * (int)gridDim.x;
    if (pair >= (p.total_tiles >> 1)) return false;
    nt = pair % p.n_tiles;
    m = (pair / p.n_tiles) * 2 + (int)rank;
  } else {
    const int t = (int)blockIdx.x + i * (int)gridDim.x;
    if (t >= p.total_tiles) return false;
    nt = t % p.n_tiles;
    m = t / p.n_tiles;
  }
  const int tiles_per_sample = p.tiles_x * p.tiles_y;
  c->sample = m / tiles_per_sample;
  m -= c->sample * tiles_per_sample;
  const int ty = m / p.tiles_x;
  c->x0 = (m - ty * p.tiles_x) * p.tw;
  c->y0 = ty * p.th;
  c->n0 = nt * block_n;
  return true;
}

template <int BLOCK_N, int BK, int NSPLIT, int CW, bool PAIR, int MT = 1, int FMT = 0, bool HALO = false>
__global__ void __launch_bounds__(kConvThreads)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmO, const ConvTcParams p, const int stages, const int nbuf, const int ngroups) {
  using L = ConvTcSmem<BLOCK_N, BK, NSPLIT, CW, PAIR, MT, FMT, HALO>;
  static_assert(!HALO || (FMT >= 2 && MT == 1 && BK == 64) ||
                    (FMT == 0 && NSPLIT == 3 && MT == 2 && !PAIR && BLOCK_N == 64 && BK == 32),
                "row-halo stages: fp16 layers (128-byte rows), or the bf16x3 N = 64 super-tile (64-byte rows)");
  static_assert(BLOCK_N % CW == 0 && (CW == 32 || CW == 64), "bad store chunk width");
  static_assert(FMT == 0 || (FMT == 1 && NSPLIT == 3) || ((FMT == 2 || FMT == 3) && NSPLIT == 1), "bad operand format");
  constexpr bool A16 = FMT >= 2;              // fp16 operands
  constexpr bool TWOK = L::kKChunks == 2;     // plane 1 of a stage = the NEXT BK channels (not the lo plane)
  constexpr bool O16 = FMT == 1 || FMT == 2;  // one fp16 output plane
  constexpr int OPL = L::kOutPlanes;
  static_assert(MT == 1 || (MT == 2 && !PAIR && 4 * BLOCK_N <= 512), "super-tiles: two pixel tiles, one CTA");
  constexpr uint32_t kRowBytes = BK * 2;  // 128 -> SW128, 64 -> SW64
  // Stacked B (N <= 64, bf16x3): one tcgen05.mma costs max(math, shared-memory operand read) and the 128-row A
  // tile alone takes 32 cycles to read, so a 128 x 64 x 16 MMA runs 48 cycles for 32 cycles of math
  // (tools/mma_bench.cu).  The hi and lo weight tiles are adjacent in shared memory, so a_hi * [w_hi ; w_lo] is ONE
  // MMA of N = 2 * BLOCK_N into a double-width accumulator (64 cycles, all math) and a_lo * w_hi adds into its
  // first half: 112 cycles per k-step instead of 144; the epilogue sums the two halves.
  constexpr bool STACK = NSPLIT == 3 && !PAIR && BLOCK_N <= 64 && 4 * MT * BLOCK_N <= 512;
  constexpr uint32_t ACCW = STACK ? 2 * BLOCK_N : BLOCK_N;  // TMEM columns of one pixel tile's accumulator
  constexpr uint32_t kTmemCols = 2 * MT * ACCW;  // two accumulator stages of MT tiles (power of two >= 64)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled tiles (checked below; the host adds no slack)
  uint8_t* smem = smem_raw;
  uint8_t* staging = smem + (size_t)stages * L::kStageBytes;  // 1024-aligned (stage bytes are a multiple of 1024)
  uint8_t* tail = staging + (size_t)ngroups * nbuf * L::kOutBufBytes;  // epilogue groups x nbuf buffers
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + 16;
  uint64_t* tfull_bar = empty_bar + 16;   // [2] accumulator ready
  uint64_t* tempty_bar = tfull_bar + 2;   // [2] accumulator drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* es = reinterpret_cast<float*>(tail + L::kTailBytes);  // [2 groups][2][BLOCK_N]
  float* eh = es + 4 * BLOCK_N;                                // [2 groups][2][BLOCK_N]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = HALO ? 3 * p.chunks : p.taps * p.chunks;  // HALO: (filter column, channel chunk)

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) {
      printf("b200al: dynamic shared memory is not 1024-byte aligned\n");
      __trap();
    }
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), PAIR ? 2 : 1);  // pair: both CTAs' producers arrive on the leader's barrier
      mbar_init(smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tfull_bar[i]), 1);
      mbar_init(smem_u32(&tempty_bar[i]), (PAIR || MT == 2) ? 8 : 4);  // one arrive per epilogue warp that reads it
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if (PAIR) { tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols); tmem_relinquish_pair(); }
    else { tmem_alloc(smem_u32(tmem_slot), kTmemCols); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  if (PAIR) cluster_sync_all();  // the peer's barriers and TMEM exist before anything is signalled at them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      TileCoord tc, tc1;
      for (int i = 0; get_tile(p, i, BLOCK_N, rank, &tc); ++i) {
        if (MT == 2) get_tile(p, i, BLOCK_N, 1u, &tc1);
        for (int kb = 0; kb < num_kb; ++kb) {
          if (HALO && NSPLIT == 3) {
            // bf16x3 super-tile: [tile 0: hi box, lo box][tile 1: hi, lo][filter row 0: w_hi, w_lo][row 1][row 2]
            const int s = kb / p.chunks, ch = kb - s * p.chunks;
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
            uint8_t* st = smem + (size_t)stage * L::kStageBytes;
            const int c = p.a_c0[s] + ch * BK, c1 = c + p.a_lo_off;
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_arrive_expect_tx(fb, L::kStageBytes);
            tma_load_5d(smem_u32(st), &tmA, fb, c, tc.x0 + p.a_dx[s], 0, tc.y0 - 1, tc.sample);
            tma_load_5d(smem_u32(st + L::kABytes), &tmA, fb, c1, tc.x0 + p.a_dx[s], 0, tc.y0 - 1, tc.sample);
            tma_load_5d(smem_u32(st + 2 * L::kABytes), &tmA, fb, c, tc1.x0 + p.a_dx[s], 0, tc1.y0 - 1, tc1.sample);
            tma_load_5d(smem_u32(st + 3 * L::kABytes), &tmA, fb, c1, tc1.x0 + p.a_dx[s], 0, tc1.y0 - 1, tc1.sample);
            for (int r = 0; r < 3; ++r)
              for (int pl = 0; pl < 2; ++pl) {
                uint8_t* bdst = st + 4 * L::kABytes + (r * 2 + pl) * L::kBBytes;
                const int row = (pl * p.taps + r * 3 + s) * p.n_total + tc.n0;
                tma_load_2d(smem_u32(bdst), &tmB, fb, ch * BK, row);
                tma_load_2d(smem_u32(bdst + L::kBBytes / 2), &tmB, fb, ch * BK, row + BLOCK_N / 2);
              }
            if (++stage == stages) { stage = 0; phase ^= 1; }
            continue;
          }
          if (HALO) {
            const int s = kb / p.chunks, ch = kb - s * p.chunks;
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
            uint8_t* st = smem + (size_t)stage * L::kStageBytes;
            const int c = p.a_c0[s] + ch * BK;
            const int cx = tc.x0 + p.a_dx[s], cy = tc.y0 - 1;  // rows y0-1 .. y0+8 (out-of-range rows arrive as zeros)
            if (PAIR) {
              const uint32_t own = smem_u32(&full_bar[stage]);
              const uint32_t fb = rank == 0 ? own : mapa_shared(own, 0);
              if (rank == 0) mbar_arrive_expect_tx(own, 2 * L::kStageBytes);
              else mbar_arrive_remote(fb);
              tma_load_5d_pair(smem_u32(st), &tmA, fb, c, cx, 0, cy, tc.sample);
              for (int r = 0; r < 3; ++r)
                tma_load_2d_pair(smem_u32(st + L::kABytes + r * L::kBBytes), &tmB, fb, ch * BK,
                                 (r * 3 + s) * p.n_total + tc.n0 + (int)rank * (BLOCK_N / 2));
            } else {
              const uint32_t fb = smem_u32(&full_bar[stage]);
              mbar_arrive_expect_tx(fb, L::kStageBytes);
              tma_load_5d(smem_u32(st), &tmA, fb, c, cx, 0, cy, tc.sample);
              for (int r = 0; r < 3; ++r) {
                uint8_t* bdst = st + L::kABytes + r * L::kBBytes;
                const int row = (r * 3 + s) * p.n_total + tc.n0;
                tma_load_2d(smem_u32(bdst), &tmB, fb, ch * BK, row);
                tma_load_2d(smem_u32(bdst + L::kBBytes / 2), &tmB, fb, ch * BK, row + BLOCK_N / 2);
              }
            }
            if (++stage == stages) { stage = 0; phase ^= 1; }
            continue;
          }
          const int tap = kb / p.chunks;
          const int ch = kb - tap * p.chunks;
          const int r = tap / 3, s = tap - r * 3;
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
          uint8_t* st = smem + (size_t)stage * L::kStageBytes;
          const int c = p.a_c0[s] + ch * BK * L::kKChunks;
          const int c1 = TWOK ? c + BK : c + p.a_lo_off;  // second plane: next channel chunk, or the lo plane
          const int cx = tc.x0 + p.a_dx[s], cy = tc.y0 + p.a_dy[r], cp = p.a_p[r];
          if (PAIR) {
            // this CTA's pixel tile and its half of the weight tile; all bytes complete on the LEADER's barrier
            const uint32_t own = smem_u32(&full_bar[stage]);
            const uint32_t fb = rank == 0 ? own : mapa_shared(own, 0);
            if (rank == 0) mbar_arrive_expect_tx(own, 2 * L::kStageBytes);
            else mbar_arrive_remote(fb);
            tma_load_5d_pair(smem_u32(st), &tmA, fb, c, cx, cp, cy, tc.sample);
            if (L::kPlanes == 2) tma_load_5d_pair(smem_u32(st + L::kABytes), &tmA, fb, c1, cx, cp, cy, tc.sample);
            for (int pl = 0; pl < L::kPlanes; ++pl)
              tma_load_2d_pair(smem_u32(st + L::kPlanes * L::kABytes + pl * L::kBBytes), &tmB, fb,
                               TWOK ? (ch * 2 + pl) * BK : ch * BK,
                               ((TWOK ? 0 : pl) * p.taps + tap) * p.n_total + tc.n0 + (int)rank * (BLOCK_N / 2));
          } else {
            const uint32_t fb = smem_u32(&full_bar[stage]);
            mbar_arrive_expect_tx(fb, L::kStageBytes);
            tma_load_5d(smem_u32(st), &tmA, fb, c, cx, cp, cy, tc.sample);
            if (L::kPlanes == 2) tma_load_5d(smem_u32(st + L::kABytes), &tmA, fb, c1, cx, cp, cy, tc.sample);
            if (MT == 2) {  // second pixel tile of the super-tile
              uint8_t* st1 = st + L::kPlanes * L::kABytes;
              const int cx1 = tc1.x0 + p.a_dx[s], cy1 = tc1.y0 + p.a_dy[r];
              tma_load_5d(smem_u32(st1), &tmA, fb, c, cx1, cp, cy1, tc1.sample);
              if (L::kPlanes == 2)
                tma_load_5d(smem_u32(st1 + L::kABytes), &tmA, fb, c1, cx1, cp, cy1, tc1.sample);
            }
            // the weight tile is fetched as two boxes of BLOCK_N/2 rows per plane (one tensor map serves both modes)
            for (int pl = 0; pl < L::kPlanes; ++pl) {
              uint8_t* bdst = st + MT * L::kPlanes * L::kABytes + pl * L::kBBytes;
              const int row = ((TWOK ? 0 : pl) * p.taps + tap) * p.n_total + tc.n0;
              const int col = TWOK ? (ch * 2 + pl) * BK : ch * BK;
              tma_load_2d(smem_u32(bdst), &tmB, fb, col, row);
              tma_load_2d(smem_u32(bdst + L::kBBytes / 2), &tmB, fb, col, row + BLOCK_N / 2);
            }
          }
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== UMMA issuer (pair: the leader CTA only) =====================
    // The whole warp walks the loop (uniform control flow keeps the descriptor arithmetic on the
    // uniform datapath); one elected lane issues the MMAs and commits.
    if (!PAIR || rank == 0) {
      constexpr uint32_t idesc = A16 ? umma_idesc_f16(PAIR ? 256 : 128, BLOCK_N) : umma_idesc_bf16(PAIR ? 256 : 128, BLOCK_N);
      constexpr uint32_t kDescHi = umma_smem_desc_hi(kRowBytes);
      const uint32_t ring_lo = umma_smem_desc_lo(smem_u32(smem));
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      TileCoord tcm;
      for (int i = 0; get_tile(p, i, BLOCK_N, rank, &tcm); ++i) {
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1);  // epilogue(s) have drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (MT * ACCW);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(smem_u32(&full_bar[stage]), phase);
          tc_fence_after();
          const uint32_t a_hi = ring_lo + (uint32_t)stage * (L::kStageBytes >> 4);
          const uint32_t a_lo = a_hi + (L::kABytes >> 4);
          const uint32_t b_hi = a_hi + ((MT * L::kPlanes * L::kABytes) >> 4);
          const uint32_t b_lo = b_hi + (L::kBBytes >> 4);
          if (HALO && NSPLIT == 3) {
            if (elect_one()) {
              constexpr uint32_t idesc2 = umma_idesc_bf16(128, 2 * BLOCK_N);
              constexpr uint32_t kA = L::kABytes >> 4;
              constexpr uint32_t kRow = (16 * kRowBytes) >> 4;   // one image row of the box (16 pixels)
              const uint32_t b0 = a_hi + 4 * kA;
#pragma unroll
              for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  const uint32_t ao = r * kRow + 2 * k;
                  const uint64_t db = make_desc(b0 + r * ((2 * L::kBBytes) >> 4) + 2 * k, kDescHi);  // [w_hi ; w_lo]
                  const uint32_t accum = (kb > 0 || r > 0 || k > 0) ? 1u : 0u;
                  umma_bf16(d_tmem, make_desc(a_hi + ao, kDescHi), db, idesc2, accum);
                  umma_bf16(d_tmem, make_desc(a_hi + kA + ao, kDescHi), db, idesc, 1u);
                  umma_bf16(d_tmem + ACCW, make_desc(a_hi + 2 * kA + ao, kDescHi), db, idesc2, accum);
                  umma_bf16(d_tmem + ACCW, make_desc(a_hi + 3 * kA + ao, kDescHi), db, idesc, 1u);
                }
              }
              umma_commit(smem_u32(&empty_bar[stage]));
              if (kb == num_kb - 1) umma_commit(smem_u32(&tfull_bar[acc]));
            }
          } else if (HALO) {
            if (elect_one()) {
              const uint32_t b0 = a_hi + (L::kABytes >> 4);
#pragma unroll
              for (int r = 0; r < 3; ++r) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  // filter row r: the 128 pixels from image row r of the box on (16 pixels x 128 B = 2048 B per row)
                  const uint64_t da = make_desc(a_hi + r * (2048 >> 4) + 2 * k, kDescHi);
                  const uint64_t db = make_desc(b0 + r * (L::kBBytes >> 4) + 2 * k, kDescHi);
                  const uint32_t accum = (kb > 0 || r > 0 || k > 0) ? 1u : 0u;
                  if (PAIR) umma_bf16_pair(d_tmem, da, db, idesc, accum);
                  else umma_bf16(d_tmem, da, db, idesc, accum);
                }
              }
              if (PAIR) {
                umma_commit_pair(smem_u32(&empty_bar[stage]));
                if (kb == num_kb - 1) umma_commit_pair(smem_u32(&tfull_bar[acc]));
              } else {
                umma_commit(smem_u32(&empty_bar[stage]));
                if (kb == num_kb - 1) umma_commit(smem_u32(&tfull_bar[acc]));
              }
            }
          } else if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_desc(a_hi + 2 * k, kDescHi);
              const uint64_t db = make_desc(b_hi + 2 * k, kDescHi);
              const uint32_t accum = (kb > 0 || k > 0) ? 1u : 0u;
              if (PAIR) {
                umma_bf16_pair(d_tmem, da, db, idesc, accum);
                if (TWOK) umma_bf16_pair(d_tmem, make_desc(a_lo + 2 * k, kDescHi), make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                if (NSPLIT == 3) {
                  umma_bf16_pair(d_tmem, da, make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                  umma_bf16_pair(d_tmem, make_desc(a_lo + 2 * k, kDescHi), db, idesc, 1u);
                }
              } else if (STACK) {
                constexpr uint32_t idesc2 = umma_idesc_bf16(128, 2 * BLOCK_N);
                umma_bf16(d_tmem, da, db, idesc2, accum);  // [hi*hi | hi*lo]: B rows run on from w_hi into w_lo
                umma_bf16(d_tmem, make_desc(a_lo + 2 * k, kDescHi), db, idesc, 1u);
                if (MT == 2) {
                  constexpr uint32_t kA1 = (L::kPlanes * L::kABytes) >> 4;
                  umma_bf16(d_tmem + ACCW, make_desc(a_hi + kA1 + 2 * k, kDescHi), db, idesc2, accum);
                  umma_bf16(d_tmem + ACCW, make_desc(a_lo + kA1 + 2 * k, kDescHi), db, idesc, 1u);
                }
              } else {
                umma_bf16(d_tmem, da, db, idesc, accum);
                if (TWOK) umma_bf16(d_tmem, make_desc(a_lo + 2 * k, kDescHi), make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                if (NSPLIT == 3) {
                  umma_bf16(d_tmem, da, make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                  umma_bf16(d_tmem, make_desc(a_lo + 2 * k, kDescHi), db, idesc, 1u);
                }
                if (MT == 2) {  // second pixel tile against the same weight tile, its own accumulator
                  constexpr uint32_t kA1 = (L::kPlanes * L::kABytes) >> 4;
                  const uint64_t da1 = make_desc(a_hi + kA1 + 2 * k, kDescHi);
                  umma_bf16(d_tmem + BLOCK_N, da1, db, idesc, accum);
                  if (TWOK)
                    umma_bf16(d_tmem + BLOCK_N, make_desc(a_lo + kA1 + 2 * k, kDescHi), make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                  if (NSPLIT == 3) {
                    umma_bf16(d_tmem + BLOCK_N, da1, make_desc(b_lo + 2 * k, kDescHi), idesc, 1u);
                    umma_bf16(d_tmem + BLOCK_N, make_desc(a_lo + kA1 + 2 * k, kDescHi), db, idesc, 1u);
                  }
                }
              }
            }
            // frees the smem slot (pair: in both CTAs) when the MMAs retire; last k-block: accumulator complete
            if (PAIR) {
              umma_commit_pair(smem_u32(&empty_bar[stage]));
              if (kb == num_kb - 1) umma_commit_pair(smem_u32(&tfull_bar[acc]));
            } else {
              umma_commit(smem_u32(&empty_bar[stage]));
              if (kb == num_kb - 1) umma_commit(smem_u32(&tfull_bar[acc]));
            }
          }
          __syncwarp();
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ===================== epilogue: two groups of 4 warps =====================
    // Group g (warps 2+4g .. 5+4g) drains accumulator stage g, i.e. every other tile of this CTA,
    // so the affine/activation/split math of two tiles runs concurrently.
    const int group = (warp - 2) >> 2;
    const int et = (threadIdx.x - 64) & 127;  // 0..127 within the group
    const int q = warp & 3;                   // TMEM lane quarter this warp may read
    int obuf = 0, ebuf = 0;
    uint8_t* gstaging = staging + (size_t)group * nbuf * L::kOutBufBytes;
    TileCoord tc;
    // MT == 2: both groups work on every super-tile, group g drains the accumulator of its g-th pixel tile
    const uint32_t trank = MT == 2 ? (uint32_t)group : rank;
    const uint32_t tcol = MT == 2 ? (uint32_t)group * ACCW : 0u;
    for (int it = 0; get_tile(p, it, BLOCK_N, trank, &tc); ++it) {
      if (MT == 1 && (group >= ngroups || (ngroups == 2 && (it & 1) != group))) continue;
      const uint32_t acc = it & 1, acc_phase = (it >> 1) & 1;
      float* esa = es + (group * 2 + ebuf) * BLOCK_N;
      float* eha = eh + (group * 2 + ebuf) * BLOCK_N;
      ebuf ^= 1;
      for (int i = et; i < BLOCK_N; i += 128) {
        const int n = tc.n0 + i;
        const int c = p.shuffle ? (n % p.cout) : n;
        float sc = p.scale[c];
        if (p.mask) sc *= p.mask[(size_t)tc.sample * p.mask_stride + c];
        esa[i] = sc;
        eha[i] = p.bias[c] * sc + p.shift[c];
      }
      if (group == 0) asm volatile("bar.sync 1, 128;" ::: "memory");
      else asm volatile("bar.sync 2, 128;" ::: "memory");
      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += CW) {
        // staging buffer `obuf` must no longer be read by an earlier TMA store of this warp
        if (elect_one()) {
          if (nbuf == 1) tma_store_wait_read<0>(); else tma_store_wait_read<1>();
        }
        __syncwarp();
        uint8_t* sub_hi = gstaging + (size_t)obuf * L::kOutBufBytes + (size_t)q * L::kOutWarpBytes;
        uint8_t* sub_lo = sub_hi + 4 * L::kOutWarpBytes;
#pragma unroll
        for (int h = 0; h < CW / 32; ++h) {
          uint32_t v[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * (MT * ACCW) + c0 + h * 32) + tcol;
          tmem_ld_32x32(taddr, v);
          if (STACK) {  // second half of the stacked accumulator: the hi*lo products
            uint32_t w[32];
            tmem_ld_32x32(taddr + BLOCK_N, w);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
          } else {
            tmem_ld_wait();
          }
          if (c0 + CW >= BLOCK_N && h == CW / 32 - 1) {  // last read of this accumulator: hand it back
            tc_fence_before();
            __syncwarp();
            if (elect_one()) {
              if (PAIR && rank != 0) mbar_arrive_remote(mapa_shared(smem_u32(&tempty_bar[acc]), 0));
              else mbar_arrive(smem_u32(&tempty_bar[acc]));
            }
          }
          uint32_t hi[16], lo[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int cc = c0 + h * 32 + 2 * j;
            const float2 sc2 = *reinterpret_cast<const float2*>(esa + cc);
            const float2 sh2 = *reinterpret_cast<const float2*>(eha + cc);
            float y0f = fmaf(__uint_as_float(v[2 * j]), sc2.x, sh2.x);
            float y1f = fmaf(__uint_as_float(v[2 * j + 1]), sc2.y, sh2.y);
            y0f = fmaxf(y0f, y0f * p.slope);  // LeakyReLU for 0 <= slope <= 1 (checked on the host)
            y1f = fmaxf(y1f, y1f * p.slope);
            if (O16) {
              const __half2 f2 = __floats2half2_rn(y0f, y1f);
              hi[j] = *reinterpret_cast<const uint32_t*>(&f2);
              continue;
            }
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(y0f, y1f);
            const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h2);
            hi[j] = hb;
            if (OPL == 2) {
              const __nv_bfloat162 l2 =
                  __floats2bfloat162_rn(y0f - __uint_as_float(hb << 16), y1f - __uint_as_float(hb & 0xffff0000u));
              lo[j] = *reinterpret_cast<const uint32_t*>(&l2);
            }
          }
          // swizzled row-major staging: row = lane, 16-byte chunk j of the row at (j ^ swz)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int chunk = h * 4 + j;
            const uint32_t off = (CW == 64) ? (uint32_t)(lane * 128 + ((chunk ^ (lane & 7)) << 4))
                                            : (uint32_t)(lane * 64 + ((chunk ^ ((lane >> 1) & 3)) << 4));
            *reinterpret_cast<uint4*>(sub_hi + off) = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
            if (OPL == 2)
              *reinterpret_cast<uint4*>(sub_lo + off) =
                  make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        int cbase = tc.n0 + c0, cx = tc.x0, cp = 0, cy = tc.y0 + 2 * q, coff = 0;
        if (p.shuffle) {  // stride-2 view of the output: (2*pitch, W/2, 2, H/2, S)
          const int quad = cbase / p.cout;
          cbase -= quad * p.cout;
          coff = (quad & 1) * (int)p.out_pitch;
          cp = quad >> 1;
        }
        if (elect_one()) {
          tma_store_5d(&tmO, smem_u32(sub_hi), coff + p.out_hi_off + cbase, cx, cp, cy, tc.sample);
          if (OPL == 2) tma_store_5d(&tmO, smem_u32(sub_lo), coff + p.out_lo_off + cbase, cx, cp, cy, tc.sample);
          tma_store_commit();
        }
        if (++obuf == nbuf) obuf = 0;
      }
    }
    if (elect_one()) tma_store_wait_all();  // all output tiles of this warp have landed
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer may still arrive on this CTA's barriers / the leader's MMAs write this TMEM
  if (warp == 2) {
    if (PAIR) tmem_dealloc_pair(tmem_base, kTmemCols);
    else tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace b200al
