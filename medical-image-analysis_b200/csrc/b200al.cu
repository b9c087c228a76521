// libb200al: C ABI + host-side plan for the acquisition pass (see include/b200al.h).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include "../../include/b200al.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "badge.cuh"
#include "common.cuh"
#include "conv_fold.cuh"
#include "conv_tc.cuh"
#include "kcenter.cuh"
#include "kcenter_fused.cuh"
#include "kmeanpp.cuh"
#include "prep.cuh"
#include "topk.cuh"
#include "uncertainty.cuh"
#include "unet_misc.cuh"

using namespace b200al;

namespace {

std::mutex g_err_mu;
std::string g_last_error;

constexpr int kChannels[5] = {32, 64, 128, 256, 512};

// ---------------------------------------------------------------------------
// activation buffers: NHWC, per pixel [hi plane (C_total) | lo plane (C_total)] bf16
// ---------------------------------------------------------------------------
enum BufId { E0A, CAT3, E1A, CAT2, E2A, CAT1, E3A, CAT0, E4A, E4B, D0A, D0B, D1A, D1B, D2A, D2B, D3A, D3B, NUM_BUFS };

struct ActBuf {
  __nv_bfloat16* ptr = nullptr;
  int level = 0;    // resolution level: H >> level
  int c_total = 0;  // channels per plane
};

struct ActView {  // a channel window of a buffer
  int buf, c_off, c;
};

enum LayerKind { KIND_CONV3 = 0, KIND_CONVT = 1 };

struct Layer {
  int kind, cin, cout, stride;
  ActView in, out;
  int block = -1;  // PlainBlock index (mask offset source) or -1
  int up = -1;     // upsample index or -1
  int mask_off = 0;
  int bn = 0, bk = 0, stages = 0, ctas_per_sm = 1, cw = 32, nbuf = 1, ngroups = 2;
  __nv_bfloat16* wp = nullptr;  // packed weights [plane][tap][N][Cin]
  __nv_bfloat16* wp2 = nullptr; // folded Cin=32 layers: [tap][N][hi 32 | lo 32]
  float *scale = nullptr, *bias = nullptr, *shift = nullptr;
  CUtensorMap tmA, tmB, tmO;
  ConvTcParams p;
  bool pair = false;   // CTA pairs running M=256 cta_group::2 MMAs (conv_tc.cuh, template PAIR)
  int mt = 1;          // 2: super-tiles of two pixel tiles sharing one weight tile per CTA (conv_tc.cuh, template MT)
  bool fold = false;   // column-folded kernel (conv_fold.cuh): stride-1 3x3, Cout = 32, Cin in {32, 64}
  ConvFoldParams fp;
  int slots = 0;
  double flops_per_sample = 0;
  int index = 0;        // position in forward order (timing records)
  // mixed-precision scoring path (conv_tc.cuh FMT): nsplit 0 = the context's; fmt 1 = fp16 single-plane output,
  // 2 = fp16 single-plane operands and output, 3 = fp16 operands, bf16 hi/lo output
  int nsplit = 0, fmt = 0;
  bool halo = false;       // fp16 stride-1 3x3: row-halo stages (conv_tc.cuh, template HALO)
  __half* wp16 = nullptr;  // packed single-plane fp16 weights [tap][N][Cin] (fmt >= 2)
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

struct b200al_ctx {
  int device = 0, K = 0, H = 0, W = 0, max_samples = 0, precision = 0;
  int nsplit = 3;
  bool weights_loaded = false;
  std::string err;
  EncodeTiledFn encode = nullptr;
  ActBuf bufs[NUM_BUFS];
  std::vector<Layer> layers;  // 21 tensor-core layers in forward order (first conv and head are separate)
  // B200AL_MIXED: the same 21 layers as the SCORING path runs them - layers [mix_first, mix_last] contract single
  // fp16 operands, layer mix_first - 1 already stores fp16 for them, mix_last stores a bf16 pair again.  The
  // embedding / BADGE / test paths keep using `layers` (bf16x3 throughout).
  std::vector<Layer> mix_layers;
  int mix_first = -1, mix_last = -1;
  // first conv + head
  float *first_w = nullptr, *first_bias = nullptr, *first_scale = nullptr, *first_shift = nullptr;
  FirstConvConsts first_consts;  // host copy of the folded first-conv constants (passed as a kernel parameter)
  float *head_w = nullptr, *head_b = nullptr;
  float* ones = nullptr;   // [512] ones  (convT scale)
  float* zeros = nullptr;  // [512] zeros (convT shift)
  float* mask = nullptr;   // [max_samples][2944]
  float* logits_buf = nullptr;  // [max_samples][K][H][W] when the head is fused into the last block
  bool fuse_head = false;
  double* partial = nullptr;
  int partial_blocks = 0;
  double* dice_partial = nullptr;  // [max_samples][kDiceBlocks][3 * kMaxClasses]
  double* badge_stats = nullptr;  // [max_samples][kBadgeBlocks][3K]
  double* badge_grad = nullptr;   // [max_samples][kBadgeBlocks][K*32]
  float* badge_coef = nullptr;    // [max_samples][2K]
  double* prep_partials = nullptr;  // [max_samples][ceil(H / kPrepBand)][2] z-score partial sums
  float slope = 0.01f;
  // k-center / top-k scratch
  float *kc_part_val = nullptr, *kc_centre = nullptr;
  int* kc_part_idx = nullptr;
  int kc_blocks = 0;
  float* kc_pp_val = nullptr;      // [2][kc_blocks] partials of the persistent k-center kernel
  long long* kc_pp_idx = nullptr;  // [2][kc_blocks]
  TopkState* tk_state = nullptr;
  // scratch of b200al_topk / b200al_kcenter_fused: sized by b200al_workspace_reserve, never inside a hot call
  char* ws = nullptr;
  size_t ws_bytes = 0;
  // bookkeeping
  long long launches = 0;
  bool timing = false;
  std::vector<cudaEvent_t> ev;
  std::vector<int> ev_layer;
  double layer_ms[32] = {0};
  size_t ev_used = 0;
  double conv_ms = 0, conv_flops = 0;
  long long conv_launches = 0, conv_timed = 0;
  std::vector<void*> allocs;
  int sm_count = 148;
};

namespace {

int set_err(b200al_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  {
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_last_error = buf;
  }
  if (ctx) ctx->err = buf;
  return code;
}

#define CUDA_TRY(ctx, expr)                                                                        \
  do {                                                                                             \
    cudaError_t e__ = (expr);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return set_err(ctx, B200AL_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), \
                     __FILE__, __LINE__);                                                          \
  } while (0)

template <typename T>
int dev_alloc(b200al_ctx* ctx, T** p, size_t count) {
  void* q = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&q, count * sizeof(T)));
  ctx->allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return 0;
}

int level_of(int buf) {
  switch (buf) {
    case E0A: case CAT3: case D3A: case D3B: return 0;
    case E1A: case CAT2: case D2A: case D2B: return 1;
    case E2A: case CAT1: case D1A: case D1B: return 2;
    case E3A: case CAT0: case D0A: case D0B: return 3;
    default: return 4;
  }
}
int ctotal_of(int buf) {
  switch (buf) {
    case E0A: case D3A: case D3B: return 32;
    case CAT3: case E1A: case D2A: case D2B: return 64;
    case CAT2: case E2A: case D1A: case D1B: return 128;
    case CAT1: case E3A: case D0A: case D0B: return 256;
    case CAT0: case E4A: case E4B: return 512;
  }
  return 0;
}

}  // namespace

// ---------------------------------------------------------------------------
// TMA descriptors
// ---------------------------------------------------------------------------
namespace {

int load_encode_fn(b200al_ctx* ctx) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CUDA_TRY(ctx, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess)
    return set_err(ctx, B200AL_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  ctx->encode = reinterpret_cast<EncodeTiledFn>(fn);
  return 0;
}

int env_int(const char* name, int dflt);

// A operand: activations [S, H, W, pitch] viewed 5-D.  stride 1: (pitch, W, 1, H, S);
// stride 2: (2*pitch, W/2, 2, H/2, S) so that every filter tap is a dense box.
int make_tma_act(b200al_ctx* ctx, CUtensorMap* tm, const __nv_bfloat16* base, int S, int H, int W, int pitch,
                 int stride, int bk, int tw, int th, int win_ch = 0) {
  cuuint64_t dims[5], strides[4];
  const cuuint64_t pb = (cuuint64_t)pitch * 2;
  if (stride == 1) {
    dims[0] = pitch; dims[1] = W; dims[2] = 1; dims[3] = H; dims[4] = S;
    strides[0] = pb; strides[1] = pb * W; strides[2] = pb * W; strides[3] = pb * W * H;
  } else {
    dims[0] = 2 * pitch; dims[1] = W / 2; dims[2] = 2; dims[3] = H / 2; dims[4] = S;
    strides[0] = 2 * pb; strides[1] = pb * W; strides[2] = 2 * pb * W; strides[3] = pb * W * H;
  }
  cuuint32_t box[5] = {(cuuint32_t)bk, (cuuint32_t)tw, 1, (cuuint32_t)th, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  // L2 promotion.  Stride-1 boxes sweep whole pixels, 256 B is free.  Stride-2 boxes skip every other pixel and
  // the hi and lo planes of one pixel arrive through two different boxes (and in a CAT buffer the other half of
  // the plane belongs to the decoder): with a promotion wider than the channel window of one plane the requests
  // fetch bytes nobody reads, and ncu shows DRAM reads at exactly 2x the tensor
  // (profiles/r01_l2promo_stride2.txt: encoder.levels.1.0 1077 -> 537 MB per 64 samples at 64 B,
  // encoder.levels.2.0 538 -> 269 MB at 128 B).  So: the window of one plane (win_ch channels), capped at 128 B.
  int promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
  if (stride == 2) {
    const int win_bytes = (win_ch > 0 ? win_ch : pitch / 2) * 2;
    promo = win_bytes <= 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    promo = env_int("B200AL_L2PROMO_S2", promo);
  }
  CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<__nv_bfloat16*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                           (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_err(ctx, B200AL_ERR_CUDA, "cuTensorMapEncodeTiled(act) failed: %d", (int)r);
  return 0;
}

// B operand: packed weights [rows = planes*taps*N][Cin], K-major.
int make_tma_wgt(b200al_ctx* ctx, CUtensorMap* tm, const __nv_bfloat16* base, int rows, int cin, int bk, int bn) {
  cuuint64_t dims[2] = {(cuuint64_t)cin, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cin * 2};
  cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)bn};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<__nv_bfloat16*>(base), dims, strides,
                           box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_err(ctx, B200AL_ERR_CUDA, "cuTensorMapEncodeTiled(wgt) failed: %d", (int)r);
  return 0;
}

// ---------------------------------------------------------------------------
// tcgen05 conv dispatch
// ---------------------------------------------------------------------------
typedef void (*ConvKernelFn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const ConvTcParams, const int,
                             const int, const int);

template <int BN, int BK, int CW>
ConvKernelFn pick_split(int nsplit) {
  return nsplit == 3 ? (ConvKernelFn)conv_tc_kernel<BN, BK, 3, CW, false> : (ConvKernelFn)conv_tc_kernel<BN, BK, 1, CW, false>;
}
template <int BN, int BK>
ConvKernelFn pick_cw(int nsplit, int cw) {
  if constexpr (BN >= 64) {
    if (cw == 64) return pick_split<BN, BK, 64>(nsplit);
  }
  return cw == 32 ? pick_split<BN, BK, 32>(nsplit) : nullptr;
}
ConvKernelFn pick_kernel(int bn, int bk, int nsplit, int cw, bool pair = false, int mt = 1) {
  if (mt == 2) {  // CW = 32, two epilogue groups
    if (pair || cw != 32) return nullptr;
#define MT2_KERNEL(BN_, BK_) \
  (nsplit == 3 ? (ConvKernelFn)conv_tc_kernel<BN_, BK_, 3, 32, false, 2> : (ConvKernelFn)conv_tc_kernel<BN_, BK_, 1, 32, false, 2>)
    if (bn == 64 && bk == 64) return MT2_KERNEL(64, 64);
    if (bn == 64 && bk == 32) return MT2_KERNEL(64, 32);
    if (bn == 128 && bk == 32) return MT2_KERNEL(128, 32);
#undef MT2_KERNEL
    return nullptr;
  }
  if (pair) {  // BK = 64, CW = 32 only
    if (bk != 64 || cw != 32) return nullptr;
    if (bn == 128) return nsplit == 3 ? (ConvKernelFn)conv_tc_kernel<128, 64, 3, 32, true> : (ConvKernelFn)conv_tc_kernel<128, 64, 1, 32, true>;
    if (bn == 256) return nsplit == 3 ? (ConvKernelFn)conv_tc_kernel<256, 64, 3, 32, true> : (ConvKernelFn)conv_tc_kernel<256, 64, 1, 32, true>;
    return nullptr;
  }
  if (bk == 32) {
    if (bn == 32) return pick_cw<32, 32>(nsplit, cw);
    if (bn == 64) return pick_cw<64, 32>(nsplit, cw);
    return nullptr;
  }
  switch (bn) {
    case 32: return pick_cw<32, 64>(nsplit, cw);
    case 64: return pick_cw<64, 64>(nsplit, cw);
    case 128: return pick_cw<128, 64>(nsplit, cw);
    case 256: return pick_cw<256, 64>(nsplit, cw);
  }
  return nullptr;
}
// pair: each CTA of the pair holds half of the weight tile's rows
// planes: operand planes per pipeline stage - 2 for the bf16 hi/lo pair and for the fp16 layers (two consecutive
// k-chunks per stage, conv_tc.cuh kKChunks), else 1
int stage_bytes(int bn, int bk, int planes, bool pair = false, int mt = 1, bool halo = false) {
  // one 10-row box (per plane and pixel tile) + three filter rows of weights
  if (halo) return planes * (mt * 160 * bk * 2 + 3 * (pair ? bn / 2 : bn) * bk * 2);
  return planes * (mt * 128 * bk * 2 + (pair ? bn / 2 : bn) * bk * 2);
}
int out_buf_bytes(int cw, int oplanes) { return oplanes * 128 * cw * 2; }
int conv_smem_bytes(int bn, int bk, int planes, int oplanes, int stages, int cw, int nbufs_total, bool pair = false,
                    int mt = 1, bool halo = false) {
  return stages * stage_bytes(bn, bk, planes, pair, mt, halo) + nbufs_total * out_buf_bytes(cw, oplanes) + 1024 +
         8 * bn * 4;
}
// per-layer operand split and number of output planes (conv_tc.cuh FMT)
int layer_nsplit(const b200al_ctx* ctx, const Layer& L);
int layer_oplanes(const b200al_ctx* ctx, const Layer& L);
int layer_planes(const b200al_ctx* ctx, const Layer& L);
ConvKernelFn pick_kernel_fmt(int bn, int bk, int cw, bool pair, int mt, int fmt, bool halo = false);
ConvKernelFn layer_kernel(const b200al_ctx* ctx, const Layer& L);
int layer_smem(const b200al_ctx* ctx, const Layer& L);

typedef void (*FoldKernelFn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const ConvFoldParams, const int);
FoldKernelFn pick_fold_kernel(int cin, int nsplit, int headk = 0) {
  if (cin == 32) {
    switch (headk) {
      case 0: return nsplit == 3 ? (FoldKernelFn)conv_fold_kernel<32, 3, 0> : (FoldKernelFn)conv_fold_kernel<32, 1, 0>;
      case 2: return nsplit == 3 ? (FoldKernelFn)conv_fold_kernel<32, 3, 2> : (FoldKernelFn)conv_fold_kernel<32, 1, 2>;
      case 3: return nsplit == 3 ? (FoldKernelFn)conv_fold_kernel<32, 3, 3> : (FoldKernelFn)conv_fold_kernel<32, 1, 3>;
      case 4: return nsplit == 3 ? (FoldKernelFn)conv_fold_kernel<32, 3, 4> : (FoldKernelFn)conv_fold_kernel<32, 1, 4>;
    }
    return nullptr;
  }
  if (cin == 64 && headk == 0)
    return nsplit == 3 ? (FoldKernelFn)conv_fold_kernel<64, 3, 0> : (FoldKernelFn)conv_fold_kernel<64, 1, 0>;
  return nullptr;
}
int fold_smem_bytes(int cin, int nsplit, int slots) {
  const int planes = nsplit == 3 ? 2 : 1;
  const int row = cin == 32 ? 128 : cin * 2;  // Cin = 32: hi and lo planes share one 128-byte row
  return (cin == 32 ? 1 : planes) * 3 * kFoldN * row + slots * 6 * 32 * row + 8 * planes * 2048 + 3072;
}

// Planner / debug switches (B200AL_VERBOSE, B200AL_MT2, B200AL_FOLD_DBG, ...) exist in TUNING builds only
// (`B200AL_TUNING_BUILD=1 python -c "import __graft_entry__ as g; g.build()"` adds -DB200AL_TUNING; tools/ab.sh and
// the experiments recorded in DESIGN.md used them).  The product library reads no environment variable.
int env_int(const char* name, int dflt) {
#ifdef B200AL_TUNING
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
#else
  (void)name;
  return dflt;
#endif
}

void choose_tile(int W, int* tw, int* th) {
  *tw = W >= 16 ? 16 : W;
  *th = 128 / *tw;
}

// Fill the parts of a layer's plan that depend only on shapes.
int plan_layer(b200al_ctx* ctx, Layer& L) {
  const ActBuf& ib = ctx->bufs[L.in.buf];
  const ActBuf& ob = ctx->bufs[L.out.buf];
  const int inH = ctx->H >> ib.level, inW = ctx->W >> ib.level;
  const int outH = ctx->H >> ob.level, outW = ctx->W >> ob.level;
  const int in_pitch = 2 * ib.c_total, out_pitch = 2 * ob.c_total;
  ConvTcParams& p = L.p;
  memset(&p, 0, sizeof(p));
  L.bk = (L.cin % 64 == 0) ? 64 : 32;
  const int n_total = L.kind == KIND_CONVT ? 4 * L.cout : L.cout;
  const int bn_max = env_int("B200AL_BN_MAX", 128);
  L.bn = n_total < bn_max ? n_total : bn_max;
  if (L.kind == KIND_CONVT && L.bn > L.cout && (L.bn % L.cout) != 0) L.bn = L.cout;
  if (L.bk == 32 && L.bn > 64) L.bn = 64;
  // The generic kernel is bound by the L2->SM operand stream (64 KB per 12 MMAs at 128x128 tiles is twice
  // what the L2 delivers per SM).  CTA pairs (cta_group::2, M = 256) split every weight tile between two
  // SMs, and N = 256 tiles halve the activation bytes per flop again.
  // Measured (profiles/r01_perlayer_v7_pair.txt): 256-wide pair tiles gain 3-15% on the Cout >= 256 layers,
  // 128-wide pair tiles LOSE 30-50% against the single-CTA 128x128 tile, so pairs are used at N >= 256 only.
  L.pair = L.bk == 64 && n_total >= env_int("B200AL_PAIR_MIN_N", 256) && env_int("B200AL_PAIR", 1) != 0;
  if (L.pair) {
    const int pbn = env_int("B200AL_PAIR_BN", 256);
    L.bn = n_total < pbn ? n_total : pbn;
    if (L.bn != 128 && L.bn != 256) L.pair = false, L.bn = n_total < bn_max ? n_total : bn_max;
  }
  // smem policy: small tiles aim for 2 CTAs/SM (<= ~110 KB each), large ones take the whole SM.
  const int ns = layer_nsplit(ctx, L), opl = layer_oplanes(ctx, L), planes = layer_planes(ctx, L);
  const int sb = stage_bytes(L.bn, L.bk, planes, L.pair);
  const int kMaxSmem = 227 * 1024;
  const int fixed = 1024 + 8 * L.bn * 4;
  L.cw = (L.bn >= 64 && L.cout >= 64) ? 64 : 32;
  L.cw = env_int("B200AL_CW", L.cw) == 32 ? 32 : L.cw;
  if (L.pair) L.cw = 32;
  // Short k-loops (full-resolution layers, transposed convs) are epilogue-heavy: two epilogue groups.
  // Long k-loops are MMA-bound: one group, the shared memory goes to pipeline stages instead.
  const int num_kb = (L.kind == KIND_CONV3 ? 9 : 1) * (L.cin / L.bk);
  L.ngroups = env_int("B200AL_NGROUPS", num_kb <= 18 ? 2 : 1);
  L.nbuf = L.ngroups == 2 ? 1 : 2;
  const int budget2 = env_int("B200AL_SMEM_BUDGET", 110 * 1024);
  int st = L.pair ? 0 : (budget2 - fixed - L.ngroups * L.nbuf * out_buf_bytes(L.cw, opl)) / sb;
  if (st < 3) {  // one CTA per SM
    st = (kMaxSmem - fixed - L.ngroups * L.nbuf * out_buf_bytes(L.cw, opl)) / sb;
    if (st < 3 && L.ngroups * L.nbuf > 1) {
      if (L.ngroups == 1) L.nbuf = 1;
      st = (kMaxSmem - fixed - L.ngroups * L.nbuf * out_buf_bytes(L.cw, opl)) / sb;
    }
    if (st < 3 && L.cw == 64) { L.cw = 32; st = (kMaxSmem - fixed - L.ngroups * L.nbuf * out_buf_bytes(32, opl)) / sb; }
  }
  if (st > 8) st = 8;
  if (st < 2) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "conv tile does not fit shared memory");
  L.stages = st;
  // Super-tiles (template MT = 2): the N <= 128 layers are bound by the L2->SM operand stream; two pixel tiles per
  // CTA share every weight tile, one CTA per SM with all of its shared memory.
  L.mt = 1;
  L.halo = false;
  // Same-box A/B of bench.py (3 alternations): 1542 -> 1580 images/s (+2.5 %) with SM clocks 1477 -> 1458 MHz under
  // the power cap; per layer: stride-2 32->64 -20 %, 128->64 at 128^2 -8 %, 128->128 at 64^2 -9 %.
  // B200AL_MT2: bit 0 = the BN = 64 layers, bit 1 = the BN = 128 layers (BK = 32 there), 0 = off.
  const int mt2 = env_int("B200AL_MT2", 3);
  // (not for Cin >= 256: those k-loops are long enough to be MMA-bound already and lose with the smaller BK)
  if (mt2 && L.kind == KIND_CONV3 && !L.pair && L.cin <= 128 && L.bn >= 64 && L.bn <= 128 &&
      (mt2 & (L.bn == 64 ? 1 : 2)) && !(L.fmt == 2 && L.bn == 128 && !env_int("B200AL_MT2_F16", 0))) {
    // (fp16 128 -> 128 layers: the single 128 x 64-channel x 2-chunk tile measured 1.46 -> 1.33 us per sample against
    // the super-tile with 32-channel chunks: longer stages matter more than the shared weight tile there)
    int bk2 = L.bk;
    if (L.bn == 128) bk2 = 32;  // 2 x 128 x 64-channel tiles + a 128-row weight tile leave room for one stage only
    if (L.bn == 64 && env_int("B200AL_MT2_BK", 64) == 32) bk2 = 32;
    const int sb2 = stage_bytes(L.bn, bk2, planes, false, 2);
    const int st2 = (kMaxSmem - fixed - 2 * out_buf_bytes(32, opl)) / sb2;
    if (st2 >= 2 && (L.fmt ? pick_kernel_fmt(L.bn, bk2, 32, false, 2, L.fmt) : pick_kernel(L.bn, bk2, ns, 32, false, 2))) {
      L.mt = 2; L.bk = bk2; L.cw = 32; L.ngroups = 2; L.nbuf = 1;
      L.stages = st2 > 8 ? 8 : st2;
    }
  }
  // The same for the bf16x3 N = 64 super-tile (decoder.levels.2.0: 128 -> 64 at 128^2), whose stages of 80 KB per 896
  // MMA cycles ask the L2 -> SM path for 89 B/clk: with 64-byte rows (BK = 32) a stage is four 10-row boxes (two pixel
  // tiles x hi, lo) + six weight tiles = 64 KB for 1344 cycles.
  if (L.fmt == 0 && ns == 3 && L.kind == KIND_CONV3 && L.stride == 1 && L.mt == 2 && L.bn == 64 && !L.pair &&
      L.cin % 32 == 0 && env_int("B200AL_HALO3", 1)) {
    const int sbh = stage_bytes(64, 32, 2, false, 2, true);
    const int sth = (kMaxSmem - fixed - 2 * out_buf_bytes(32, opl)) / sbh;
    if (sth >= 3) {
      L.halo = true; L.bk = 32; L.cw = 32; L.ngroups = 2; L.nbuf = 1;
      L.stages = sth > 8 ? 8 : sth;
    }
  }
  // Row-halo stages (template HALO) for the fp16 stride-1 3x3 layers: with single operands the shared-memory refill
  // sets the pace; one 10-row activation box per filter column serves the three filter rows, so
  // a stage is 20 KB of activations + 3 weight tiles for 12 MMAs instead of 2 x 16 KB + 2 weight tiles for 8.
  if (L.fmt == 2 && L.kind == KIND_CONV3 && L.stride == 1 && L.mt == 1 && L.bk == 64 && env_int("B200AL_HALO", 1)) {
    int nb = L.nbuf;
    const int sbh = stage_bytes(L.bn, 64, 1, L.pair, 1, true);  // (planes = 1: a single chunk per stage)
    int sth = (kMaxSmem - fixed - L.ngroups * nb * out_buf_bytes(32, opl)) / sbh;
    if (sth < 3 && L.ngroups == 1 && nb == 2) {
      nb = 1;
      sth = (kMaxSmem - fixed - L.ngroups * nb * out_buf_bytes(32, opl)) / sbh;
    }
    if (sth >= 3 && pick_kernel_fmt(L.bn, 64, 32, L.pair, 1, 2, true)) {
      L.halo = true; L.cw = 32; L.nbuf = nb;
      L.stages = sth > 8 ? 8 : sth;
    }
  }
  p.taps = L.kind == KIND_CONV3 ? 9 : 1;
  // fp16 layers: two BK-channel chunks per pipeline stage (row-halo stages: one chunk, three filter rows)
  p.chunks = L.cin / (L.bk * ((L.fmt >= 2 && !L.halo) ? 2 : 1));
  if (L.fmt >= 2 && !L.halo && L.cin % (2 * L.bk))
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "fp16 layer needs Cin %% %d == 0", 2 * L.bk);
  p.n_total = n_total;
  p.a_lo_off = ib.c_total;
  const int gridW = L.kind == KIND_CONV3 ? outW : inW, gridH = L.kind == KIND_CONV3 ? outH : inH;
  choose_tile(gridW, &p.tw, &p.th);
  if (p.tw != 16 || p.th != 8)
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "image width %d too small for the 8x16 output tile", ctx->W);
  if (gridW % p.tw || gridH % p.th)
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "image size %dx%d not tileable at level %d", ctx->H, ctx->W, ob.level);
  p.tiles_x = gridW / p.tw;
  p.tiles_y = gridH / p.th;
  for (int i = 0; i < 3; ++i) {
    if (L.kind == KIND_CONVT) {
      p.a_c0[i] = L.in.c_off; p.a_dx[i] = 0; p.a_p[i] = 0; p.a_dy[i] = 0;
    } else if (L.stride == 1) {
      p.a_c0[i] = L.in.c_off; p.a_dx[i] = i - 1; p.a_p[i] = 0; p.a_dy[i] = i - 1;
    } else {
      // input pixel 2*o + i - 1: i=0 -> (o-1, parity 1); i=1 -> (o, 0); i=2 -> (o, 1)
      const int par = (i == 1) ? 0 : 1;
      p.a_c0[i] = par * in_pitch + L.in.c_off;
      p.a_dx[i] = (i == 0) ? -1 : 0;
      p.a_p[i] = par;
      p.a_dy[i] = (i == 0) ? -1 : 0;
    }
  }
  // Cout = 64 runs as two launches of 32 output channels each (the resident weights of one half fit, all 64 do not)
  L.fold = L.fmt == 0 && L.kind == KIND_CONV3 && L.stride == 1 && outH % 4 == 0 && env_int("B200AL_FOLD", 1) != 0 &&
           ((L.cin == 32 && L.cout == 32 && ib.c_total == 32 && L.in.c_off == 0) ||
            (L.cin == 64 && (L.cout == 32 || (L.cout == 64 && env_int("B200AL_FOLD64", 1)))));
  if (L.fold) {
    ConvFoldParams& f = L.fp;
    memset(&f, 0, sizeof(f));
    f.a_c0 = L.in.c_off;
    f.a_lo_off = ib.c_total;
    f.tiles_x = (outW + kFoldValid - 1) / kFoldValid;
    f.tiles_y = outH / 4;
    f.slope = ctx->slope;
    f.out_hi_off = L.out.c_off;
    f.out_lo_off = ob.c_total + L.out.c_off;
    f.mask_stride = B200AL_MASK_CHANNELS;
    f.dbg = env_int("B200AL_FOLD_DBG", 0);
    int sl = 16;
    while (sl > 2 && fold_smem_bytes(L.cin, ctx->nsplit, sl) > kMaxSmem) --sl;
    sl = env_int("B200AL_FOLD_SLOTS", sl);
    if (sl < 2 || sl > 16 || fold_smem_bytes(L.cin, ctx->nsplit, sl) > kMaxSmem)
      return set_err(ctx, B200AL_ERR_UNSUPPORTED, "folded conv tile does not fit shared memory");
    L.slots = sl;
  }
  p.slope = L.kind == KIND_CONV3 ? ctx->slope : 1.0f;
  p.out = ob.ptr;
  p.out_pitch = out_pitch;
  p.out_hi_off = L.out.c_off;
  p.out_lo_off = ob.c_total + L.out.c_off;
  p.out_W = outW;
  p.out_H = outH;
  p.shuffle = L.kind == KIND_CONVT;
  p.cout = L.cout;
  p.write_lo = opl == 2;
  p.mask_stride = B200AL_MASK_CHANNELS;
  const double px = L.kind == KIND_CONV3 ? (double)outH * outW : (double)inH * inW;
  L.flops_per_sample = 2.0 * px * n_total * (double)(p.taps * L.cin);
  return 0;
}

// TMA maps of a layer for given input / output base pointers (the context's own buffers, or a test's).
int make_layer_maps(b200al_ctx* ctx, const Layer& L, const __nv_bfloat16* in, const __nv_bfloat16* out, int S,
                    CUtensorMap* tmA, CUtensorMap* tmO) {
  const ActBuf& ib = ctx->bufs[L.in.buf];
  const ActBuf& ob = ctx->bufs[L.out.buf];
  const int inH = ctx->H >> ib.level, inW = ctx->W >> ib.level;
  const int outH = ctx->H >> ob.level, outW = ctx->W >> ob.level;
  int rc;
  if (L.fold) {
    // A: 6-row x 32-column halo box, all Cin channels of one plane; O: one image row of 30 pixels x 32 channels
    //    (Cin = 32: both planes of the pixel in one 64-element row)
    if ((rc = make_tma_act(ctx, tmA, in, S, inH, inW, 2 * ib.c_total, 1, 64, 32, 6))) return rc;
    return make_tma_act(ctx, tmO, out, S, outH, outW, 2 * ob.c_total, 1, 32, kFoldValid, 1);
  }
  if ((rc = make_tma_act(ctx, tmA, in, S, inH, inW, 2 * ib.c_total, L.kind == KIND_CONV3 ? L.stride : 1, L.bk, L.p.tw,
                         L.halo ? L.p.th + 2 : L.p.th, L.cin)))
    return rc;
  // output map: box = one epilogue warp's 2 image rows x 16 px x CW channels; transposed conv uses the stride-2 view
  return make_tma_act(ctx, tmO, out, S, outH, outW, 2 * ob.c_total, L.kind == KIND_CONVT ? 2 : 1, L.cw, 16, 2);
}

int launch_conv(b200al_ctx* ctx, Layer& L, const CUtensorMap& tmA, const CUtensorMap& tmO, const ConvTcParams& p, int S,
                cudaStream_t st, float* head_logits = nullptr) {
  if (L.fold) {
    FoldKernelFn ffn = pick_fold_kernel(L.cin, ctx->nsplit, head_logits ? ctx->K : 0);
    if (!ffn) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "no folded kernel for Cin=%d head K=%d", L.cin, ctx->K);
    ConvFoldParams f = L.fp;
    f.head_w = ctx->head_w; f.head_b = ctx->head_b; f.logits = head_logits; f.H = ctx->H; f.W = ctx->W;
    f.scale = p.scale; f.bias = p.bias; f.shift = p.shift; f.mask = p.mask; f.slope = p.slope;
    f.total_tiles = S * f.tiles_x * f.tiles_y;
    dim3 fgrid((unsigned)(f.total_tiles < ctx->sm_count ? f.total_tiles : ctx->sm_count));
    const int fsmem = fold_smem_bytes(L.cin, ctx->nsplit, L.slots);
    const bool ft = ctx->timing && ctx->ev_used + 2 <= ctx->ev.size();
    if (ft) { ctx->ev_layer[ctx->ev_used / 2] = L.index; cudaEventRecord(ctx->ev[ctx->ev_used++], st); ctx->conv_flops += L.flops_per_sample * S; ctx->conv_timed++; }
    f.w_cout_total = L.cout;
    for (int co0 = 0; co0 < L.cout; co0 += 32) {
      ConvFoldParams g = f;
      g.w_co0 = co0;
      g.scale += co0; g.bias += co0; g.shift += co0;
      if (g.mask) g.mask += co0;
      g.out_hi_off += co0; g.out_lo_off += co0;
      ffn<<<fgrid, kFoldThreads, fsmem, st>>>(tmA, L.tmB, tmO, g, L.slots);
      ctx->launches++;
      ctx->conv_launches++;
    }
    if (ft) cudaEventRecord(ctx->ev[ctx->ev_used++], st);
    CUDA_TRY(ctx, cudaGetLastError());
    return 0;
  }
  ConvKernelFn fn = layer_kernel(ctx, L);
  if (!fn) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "no conv kernel for BN=%d BK=%d fmt=%d", L.bn, L.bk, L.fmt);
  ConvTcParams q = p;
  q.n_tiles = p.n_total / L.bn;
  q.total_tiles = S * p.tiles_x * p.tiles_y * q.n_tiles;
  const int resident = ctx->sm_count * L.ctas_per_sm;
  // CTA pairs with multicast weight tiles: pixel tiles come in pairs, one CTA per SM
  const int m_tiles = S * p.tiles_x * p.tiles_y;
  if ((L.pair || L.mt == 2) && m_tiles % 2)
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "odd number of pixel tiles in a paired layer");
  q.cluster = L.pair ? 1 : 0;
  q.mt = L.mt;
  const int work = L.mt == 2 ? q.total_tiles / 2 : q.total_tiles;
  unsigned nblk = (unsigned)(work < resident ? work : resident);
  if (q.cluster) nblk &= ~1u;
  const int smem = layer_smem(ctx, L);
  const bool t = ctx->timing && ctx->ev_used + 2 <= ctx->ev.size();
  if (t) { ctx->ev_layer[ctx->ev_used / 2] = L.index; cudaEventRecord(ctx->ev[ctx->ev_used++], st); ctx->conv_flops += L.flops_per_sample * S; ctx->conv_timed++; }
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nblk);
    cfg.blockDim = dim3(kConvThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = q.cluster ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_TRY(ctx, cudaLaunchKernelEx(&cfg, fn, tmA, L.tmB, tmO, (const ConvTcParams)q, (const int)L.stages, (const int)L.nbuf,
                                     (const int)L.ngroups));
  }
  if (t) cudaEventRecord(ctx->ev[ctx->ev_used++], st);
  ctx->launches++;
  ctx->conv_launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return 0;
}

}  // namespace

namespace {
int layer_nsplit(const b200al_ctx* ctx, const Layer& L) { return L.nsplit ? L.nsplit : ctx->nsplit; }
int layer_oplanes(const b200al_ctx* ctx, const Layer& L) {
  if (L.fmt == 0) return layer_nsplit(ctx, L) == 3 ? 2 : 1;
  return L.fmt == 3 ? 2 : 1;
}
int layer_planes(const b200al_ctx* ctx, const Layer& L) {
  if (L.halo && L.fmt >= 2) return 1;  // row-halo fp16 stages hold one channel chunk
  return (layer_nsplit(ctx, L) == 3 || L.fmt >= 2) ? 2 : 1;
}
// The format variants are instantiated for the tile shapes the planner gives the layers of the fp16 groups only.
ConvKernelFn pick_kernel_fmt(int bn, int bk, int cw, bool pair, int mt, int fmt, bool halo) {
  if (halo) {  // fp16 operands and output, row-halo stages (stride-1 3x3)
    if (fmt != 2 || mt != 1 || bk != 64 || cw != 32) return nullptr;
    if (pair && bn == 256) return (ConvKernelFn)conv_tc_kernel<256, 64, 1, 32, true, 1, 2, true>;
    if (!pair && bn == 128) return (ConvKernelFn)conv_tc_kernel<128, 64, 1, 32, false, 1, 2, true>;
    return nullptr;
  }
  if (fmt == 1) {  // bf16x3 operands, fp16 output: the layer in front of a group
    if (!pair && mt == 2 && bn == 128 && bk == 32 && cw == 32) return (ConvKernelFn)conv_tc_kernel<128, 32, 3, 32, false, 2, 1>;
    if (pair && bn == 256 && bk == 64 && cw == 32) return (ConvKernelFn)conv_tc_kernel<256, 64, 3, 32, true, 1, 1>;
    return nullptr;
  }
  if (fmt == 2) {  // fp16 operands and output
    if (pair && bn == 256 && bk == 64 && cw == 32) return (ConvKernelFn)conv_tc_kernel<256, 64, 1, 32, true, 1, 2>;
    if (!pair && mt == 2 && bn == 128 && bk == 32 && cw == 32) return (ConvKernelFn)conv_tc_kernel<128, 32, 1, 32, false, 2, 2>;
    if (!pair && mt == 1 && bn == 128 && bk == 64 && cw == 64) return (ConvKernelFn)conv_tc_kernel<128, 64, 1, 64, false, 1, 2>;
    if (!pair && mt == 1 && bn == 128 && bk == 64 && cw == 32) return (ConvKernelFn)conv_tc_kernel<128, 64, 1, 32, false, 1, 2>;
    return nullptr;
  }
  if (fmt == 3) {  // fp16 operands, bf16 hi/lo output: the last layer of a group (a transposed conv)
    if (pair && bn == 256 && bk == 64 && cw == 32) return (ConvKernelFn)conv_tc_kernel<256, 64, 1, 32, true, 1, 3>;
    return nullptr;
  }
  return nullptr;
}
ConvKernelFn layer_kernel(const b200al_ctx* ctx, const Layer& L) {
  if (L.fmt) return pick_kernel_fmt(L.bn, L.bk, L.cw, L.pair, L.mt, L.fmt, L.halo);
  if (L.halo) {  // bf16x3 N = 64 super-tile with row-halo stages
    if (layer_nsplit(ctx, L) == 3 && L.bn == 64 && L.bk == 32 && L.cw == 32 && !L.pair && L.mt == 2)
      return (ConvKernelFn)conv_tc_kernel<64, 32, 3, 32, false, 2, 0, true>;
    return nullptr;
  }
  return pick_kernel(L.bn, L.bk, layer_nsplit(ctx, L), L.cw, L.pair, L.mt);
}
int layer_smem(const b200al_ctx* ctx, const Layer& L) {
  return conv_smem_bytes(L.bn, L.bk, layer_planes(ctx, L), layer_oplanes(ctx, L), L.stages, L.cw, L.nbuf * L.ngroups,
                         L.pair, L.mt, L.halo);
}
}  // namespace

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
namespace {

void add_conv(b200al_ctx* ctx, int block, int cin, int cout, int stride, ActView in, ActView out, int mask_off) {
  Layer L;
  L.kind = KIND_CONV3; L.cin = cin; L.cout = cout; L.stride = stride; L.in = in; L.out = out;
  L.block = block; L.mask_off = mask_off;
  L.index = (int)ctx->layers.size();
  ctx->layers.push_back(L);
}
void add_convT(b200al_ctx* ctx, int up, int cin, int cout, ActView in, ActView out) {
  Layer L;
  L.kind = KIND_CONVT; L.cin = cin; L.cout = cout; L.stride = 1; L.in = in; L.out = out; L.up = up;
  L.index = (int)ctx->layers.size();
  ctx->layers.push_back(L);
}

// mask channel offset of PlainBlock b (forward order; sum = 2944)
int mask_offset(int b) {
  static const int c[18] = {32, 32, 64, 64, 128, 128, 256, 256, 512, 512, 256, 256, 128, 128, 64, 64, 32, 32};
  int off = 0;
  for (int i = 0; i < b; ++i) off += c[i];
  return off;
}

void build_layer_list(b200al_ctx* ctx) {
  // encoder (unet.py:53-78): block 0 (first conv) is handled by conv_first_kernel
  add_conv(ctx, 1, 32, 32, 1, {E0A, 0, 32}, {CAT3, 0, 32}, mask_offset(1));
  add_conv(ctx, 2, 32, 64, 2, {CAT3, 0, 32}, {E1A, 0, 64}, mask_offset(2));
  add_conv(ctx, 3, 64, 64, 1, {E1A, 0, 64}, {CAT2, 0, 64}, mask_offset(3));
  add_conv(ctx, 4, 64, 128, 2, {CAT2, 0, 64}, {E2A, 0, 128}, mask_offset(4));
  add_conv(ctx, 5, 128, 128, 1, {E2A, 0, 128}, {CAT1, 0, 128}, mask_offset(5));
  add_conv(ctx, 6, 128, 256, 2, {CAT1, 0, 128}, {E3A, 0, 256}, mask_offset(6));
  add_conv(ctx, 7, 256, 256, 1, {E3A, 0, 256}, {CAT0, 0, 256}, mask_offset(7));
  add_conv(ctx, 8, 256, 512, 2, {CAT0, 0, 256}, {E4A, 0, 512}, mask_offset(8));
  add_conv(ctx, 9, 512, 512, 1, {E4A, 0, 512}, {E4B, 0, 512}, mask_offset(9));
  // decoder (unet.py:208-224): upsample -> cat([skip, up]) -> two blocks
  add_convT(ctx, 0, 512, 256, {E4B, 0, 512}, {CAT0, 256, 256});
  add_conv(ctx, 10, 512, 256, 1, {CAT0, 0, 512}, {D0A, 0, 256}, mask_offset(10));
  add_conv(ctx, 11, 256, 256, 1, {D0A, 0, 256}, {D0B, 0, 256}, mask_offset(11));
  add_convT(ctx, 1, 256, 128, {D0B, 0, 256}, {CAT1, 128, 128});
  add_conv(ctx, 12, 256, 128, 1, {CAT1, 0, 256}, {D1A, 0, 128}, mask_offset(12));
  add_conv(ctx, 13, 128, 128, 1, {D1A, 0, 128}, {D1B, 0, 128}, mask_offset(13));
  add_convT(ctx, 2, 128, 64, {D1B, 0, 128}, {CAT2, 64, 64});
  add_conv(ctx, 14, 128, 64, 1, {CAT2, 0, 128}, {D2A, 0, 64}, mask_offset(14));
  add_conv(ctx, 15, 64, 64, 1, {D2A, 0, 64}, {D2B, 0, 64}, mask_offset(15));
  add_convT(ctx, 3, 64, 32, {D2B, 0, 64}, {CAT3, 32, 32});
  add_conv(ctx, 16, 64, 32, 1, {CAT3, 0, 64}, {D3A, 0, 32}, mask_offset(16));
  add_conv(ctx, 17, 32, 32, 1, {D3A, 0, 32}, {D3B, 0, 32}, mask_offset(17));
}

constexpr int kEncoderLayers = 9;  // layers[0..8] end at the bottleneck (E4B)

}  // namespace

namespace {
// Weight map, kernel attributes and residency of a layer that runs on the generic kernel (either layer list).
int finish_generic_layer(b200al_ctx* ctx, Layer& L) {
  int rc;
  if (L.fmt >= 2) {  // single fp16 plane: rows [taps][N]
    if ((rc = make_tma_wgt(ctx, &L.tmB, reinterpret_cast<const __nv_bfloat16*>(L.wp16), L.p.taps * L.p.n_total, L.cin,
                           L.bk, L.bn / 2)))
      return rc;
  } else {
    // rows: conv [2 planes][9 taps][Cout]; convT [2 planes][1][4*Cout]
    // (boxes of BN/2 rows: a weight tile is fetched as two halves, see conv_tc.cuh)
    if ((rc = make_tma_wgt(ctx, &L.tmB, L.wp, 2 * L.p.taps * L.p.n_total, L.cin, L.bk, L.bn / 2))) return rc;
  }
  ConvKernelFn fn = layer_kernel(ctx, L);
  if (!fn)
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "no kernel for BN=%d BK=%d CW=%d pair=%d mt=%d fmt=%d (layer %d)", L.bn,
                   L.bk, L.cw, (int)L.pair, L.mt, L.fmt, L.index);
  const int smem_l = layer_smem(ctx, L);
  CUDA_TRY(ctx, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  int occ = 0;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kConvThreads, smem_l));
  const int tmem_limit = 512 / (2 * L.mt * L.bn);  // two accumulator stages per CTA
  L.ctas_per_sm = occ < 1 ? 1 : (occ > tmem_limit ? tmem_limit : occ);
  if (L.ctas_per_sm > 2) L.ctas_per_sm = 2;
  if (L.pair || L.mt == 2) L.ctas_per_sm = 1;
  if (env_int("B200AL_VERBOSE", 0))
    fprintf(stderr, "b200al plan: layer %d kind=%d %d->%d s%d fmt=%d BN=%d BK=%d stages=%d CW=%d nbuf=%d groups=%d smem=%d ctas/sm=%d pair=%d mt=%d halo=%d\n",
            L.index, L.kind, L.cin, L.cout, L.stride, L.fmt, L.bn, L.bk, L.stages, L.cw, L.nbuf, L.ngroups, smem_l,
            L.ctas_per_sm, (int)L.pair, L.mt, (int)L.halo);
  return 0;
}

// B200AL_MIXED / B200AL_MIXED_DEEP: the scoring path's own copy of the layer list with an fp16 group in the deep part
// of the network (where single fp16 operands cost <= 4e-3 of logit error per layer, profiles/r02_precision.json).
int build_mix_layers(b200al_ctx* ctx, int entry, int first, int last) {
  ctx->mix_layers = ctx->layers;  // same buffers, maps, affine tables; bf16x3 plans outside the group
  ctx->mix_first = first;
  ctx->mix_last = last;
  for (int i = entry; i <= last; ++i) {
    Layer& L = ctx->mix_layers[i];
    if (L.fold) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "layer %d of the fp16 group runs on the folded kernel", i);
    L.fmt = i == entry ? 1 : (i == last ? 3 : 2);
    L.nsplit = i == entry ? 3 : 1;
    int rc;
    if ((rc = plan_layer(ctx, L))) return rc;
    if (L.fold) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "layer %d of the fp16 group was planned as folded", i);
    if (L.fmt >= 2) {
      const int taps = L.kind == KIND_CONV3 ? 9 : 4;
      if ((rc = dev_alloc(ctx, &L.wp16, (size_t)taps * L.cout * L.cin))) return rc;
    }
    if ((rc = make_layer_maps(ctx, L, ctx->bufs[L.in.buf].ptr, ctx->bufs[L.out.buf].ptr, ctx->max_samples, &L.tmA, &L.tmO)))
      return rc;
    if ((rc = finish_generic_layer(ctx, L))) return rc;
  }
  return 0;
}
}  // namespace

extern "C" int b200al_ctx_create(b200al_ctx** out, int device, int num_out, int in_channels, int H, int W,
                                 int max_samples, int precision) {
  if (!out) return set_err(nullptr, B200AL_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (in_channels != 1)
    return set_err(nullptr, B200AL_ERR_UNSUPPORTED, "in_channels=%d: only grayscale input (1) is supported", in_channels);
  if (num_out < 2 || num_out > kMaxClasses)
    return set_err(nullptr, B200AL_ERR_UNSUPPORTED, "num_out=%d outside [2,%d]", num_out, kMaxClasses);
  if (H < 128 || W < 256 || H % 128 || W % 256)
    return set_err(nullptr, B200AL_ERR_UNSUPPORTED, "H must be a multiple of 128 and W of 256 (got %dx%d)", H, W);
  if (max_samples < 1) return set_err(nullptr, B200AL_ERR_ARG, "max_samples must be >= 1");
  if (precision < B200AL_BF16X3 || precision > B200AL_MIXED_DEEP) return set_err(nullptr, B200AL_ERR_ARG, "bad precision");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
    return set_err(nullptr, B200AL_ERR_CUDA, "no CUDA device visible: libb200al has no CPU fallback");
  if (device < 0 || device >= ndev) return set_err(nullptr, B200AL_ERR_ARG, "device %d out of range", device);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10)
    return set_err(nullptr, B200AL_ERR_CUDA, "device %d is not sm_100 (compute capability %d.%d)", device, prop.major,
                   prop.minor);
  b200al_ctx* ctx = new b200al_ctx();
  ctx->device = device; ctx->K = num_out; ctx->H = H; ctx->W = W; ctx->max_samples = max_samples;
  ctx->precision = precision; ctx->nsplit = precision == B200AL_BF16 ? 1 : 3;
  ctx->sm_count = prop.multiProcessorCount;
  int rc = 0;
#define TRY_RC(x) do { rc = (x); if (rc) { g_last_error = ctx->err; b200al_ctx_destroy(ctx); return rc; } } while (0)
#define TRY_CU(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { rc = set_err(ctx, B200AL_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e__)); b200al_ctx_destroy(ctx); return rc; } } while (0)
  TRY_CU(cudaSetDevice(device));
  TRY_RC(load_encode_fn(ctx));
  for (int b = 0; b < NUM_BUFS; ++b) {
    ctx->bufs[b].level = level_of(b);
    ctx->bufs[b].c_total = ctotal_of(b);
    const size_t elems = (size_t)max_samples * (H >> level_of(b)) * (W >> level_of(b)) * 2 * ctotal_of(b);
    TRY_RC(dev_alloc(ctx, &ctx->bufs[b].ptr, elems));
  }
  build_layer_list(ctx);
  for (Layer& L : ctx->layers) {
    TRY_RC(plan_layer(ctx, L));
    const int taps = L.kind == KIND_CONV3 ? 9 : 4;
    TRY_RC(dev_alloc(ctx, &L.wp, (size_t)2 * taps * L.cout * L.cin));
    TRY_RC(dev_alloc(ctx, &L.scale, (size_t)L.cout));
    TRY_RC(dev_alloc(ctx, &L.bias, (size_t)L.cout));
    TRY_RC(dev_alloc(ctx, &L.shift, (size_t)L.cout));
    TRY_RC(make_layer_maps(ctx, L, ctx->bufs[L.in.buf].ptr, ctx->bufs[L.out.buf].ptr, max_samples, &L.tmA, &L.tmO));
    if (L.fold) {
      // weights: rows [2 planes][9 taps][32], box = the three column taps of one filter row (96 rows) x Cin
      if (L.cin == 32) {
        TRY_RC(dev_alloc(ctx, &L.wp2, (size_t)9 * L.cout * 64));
        TRY_RC(make_tma_wgt(ctx, &L.tmB, L.wp2, 9 * L.cout, 64, 64, kFoldN));
      } else {
        TRY_RC(make_tma_wgt(ctx, &L.tmB, L.wp, 2 * 9 * L.cout, L.cin, L.cin, 32));
      }
      FoldKernelFn ffn = pick_fold_kernel(L.cin, ctx->nsplit);
      TRY_CU(cudaFuncSetAttribute(ffn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
      if (&L == &ctx->layers.back() && pick_fold_kernel(L.cin, ctx->nsplit, ctx->K) && env_int("B200AL_FUSE_HEAD", 1)) {
        TRY_CU(cudaFuncSetAttribute(pick_fold_kernel(L.cin, ctx->nsplit, ctx->K),
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        TRY_RC(dev_alloc(ctx, &ctx->logits_buf, (size_t)max_samples * ctx->K * H * W));
        ctx->fuse_head = true;
      }
      if (env_int("B200AL_VERBOSE", 0))
        fprintf(stderr, "b200al plan: fold %d->%d slots=%d smem=%d\n", L.cin, L.cout, L.slots,
                fold_smem_bytes(L.cin, ctx->nsplit, L.slots));
      continue;
    }
    TRY_RC(finish_generic_layer(ctx, L));
  }
  if (precision == B200AL_MIXED) TRY_RC(build_mix_layers(ctx, 3, 4, 15));       // enc2.0 | enc2.1 .. up2
  if (precision == B200AL_MIXED_DEEP) TRY_RC(build_mix_layers(ctx, 5, 6, 12));  // enc3.0 | enc3.1 .. up1
  TRY_RC(dev_alloc(ctx, &ctx->first_w, 32 * 9));
  TRY_RC(dev_alloc(ctx, &ctx->first_bias, 32));
  TRY_RC(dev_alloc(ctx, &ctx->first_scale, 32));
  TRY_RC(dev_alloc(ctx, &ctx->first_shift, 32));
  TRY_RC(dev_alloc(ctx, &ctx->head_w, (size_t)kMaxClasses * 32));
  TRY_RC(dev_alloc(ctx, &ctx->head_b, (size_t)kMaxClasses));
  TRY_RC(dev_alloc(ctx, &ctx->ones, 512));
  TRY_RC(dev_alloc(ctx, &ctx->zeros, 512));
  TRY_RC(dev_alloc(ctx, &ctx->mask, (size_t)max_samples * B200AL_MASK_CHANNELS));
  ctx->partial_blocks = 64;
  TRY_RC(dev_alloc(ctx, &ctx->partial, (size_t)max_samples * ctx->partial_blocks));
  TRY_RC(dev_alloc(ctx, &ctx->dice_partial, (size_t)max_samples * kDiceBlocks * 3 * kMaxClasses));
  TRY_RC(dev_alloc(ctx, &ctx->prep_partials, (size_t)max_samples * ((ctx->H + kPrepBand - 1) / kPrepBand) * 2));
  ctx->kc_blocks = ctx->sm_count * 4;
  TRY_RC(dev_alloc(ctx, &ctx->kc_part_val, (size_t)ctx->kc_blocks));
  TRY_RC(dev_alloc(ctx, &ctx->kc_part_idx, (size_t)ctx->kc_blocks));
  TRY_RC(dev_alloc(ctx, &ctx->kc_centre, 2048));
  TRY_RC(dev_alloc(ctx, &ctx->kc_pp_val, (size_t)2 * ctx->kc_blocks));
  TRY_RC(dev_alloc(ctx, &ctx->kc_pp_idx, (size_t)2 * ctx->kc_blocks));
  TRY_RC(dev_alloc(ctx, &ctx->tk_state, 1));
  fill_kernel<<<2, 256>>>(ctx->ones, 1.0f, 512);
  fill_kernel<<<2, 256>>>(ctx->zeros, 0.0f, 512);
  TRY_CU(cudaDeviceSynchronize());
  ctx->ev.resize(2 * 32768);
  ctx->ev_layer.resize(32768);
  for (auto& e : ctx->ev) TRY_CU(cudaEventCreate(&e));
#undef TRY_RC
#undef TRY_CU
  *out = ctx;
  return B200AL_OK;
}

extern "C" void b200al_ctx_destroy(b200al_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  for (void* p : ctx->allocs) cudaFree(p);
  if (ctx->ws) cudaFree(ctx->ws);
  for (auto& e : ctx->ev)
    if (e) cudaEventDestroy(e);
  delete ctx;
}

extern "C" const char* b200al_last_error(const b200al_ctx* ctx) {
  if (ctx) return ctx->err.c_str();
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(g_err_mu);
  copy = g_last_error;
  return copy.c_str();
}

// ---------------------------------------------------------------------------
// weights
// ---------------------------------------------------------------------------
extern "C" int b200al_load_weights(b200al_ctx* ctx, const b200al_weights* w, void* stream) {
  if (!ctx || !w) return set_err(ctx, B200AL_ERR_ARG, "NULL argument");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (!(w->leaky_slope >= 0.f && w->leaky_slope <= 1.f))
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "LeakyReLU slope %g outside [0,1]", (double)w->leaky_slope);
  ctx->slope = w->leaky_slope;
  const b200al_block_weights& b0 = w->blocks[0];
  if (!b0.d_conv_weight || !w->d_head_weight || !w->d_head_bias) return set_err(ctx, B200AL_ERR_ARG, "NULL weight pointer");
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_w, b0.d_conv_weight, 32 * 9 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_bias, b0.d_conv_bias, 32 * sizeof(float), cudaMemcpyDeviceToDevice, st));
  fold_bn_kernel<<<1, 32, 0, st>>>(b0.d_bn_gamma, b0.d_bn_beta, b0.d_bn_mean, b0.d_bn_var, w->bn_eps, ctx->first_scale,
                                   ctx->first_shift, 32);
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_consts.w, ctx->first_w, sizeof(ctx->first_consts.w), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_consts.bias, ctx->first_bias, 32 * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_consts.scale, ctx->first_scale, 32 * sizeof(float), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->first_consts.shift, ctx->first_shift, 32 * sizeof(float), cudaMemcpyDeviceToHost, st));
  for (Layer& L : ctx->layers) {
    L.p.slope = L.kind == KIND_CONV3 ? ctx->slope : 1.0f;
    if (L.kind == KIND_CONV3) {
      const b200al_block_weights& b = w->blocks[L.block];
      if (!b.d_conv_weight || !b.d_conv_bias || !b.d_bn_gamma || !b.d_bn_beta || !b.d_bn_mean || !b.d_bn_var)
        return set_err(ctx, B200AL_ERR_ARG, "NULL weight pointer in block %d", L.block);
      pack_conv3x3_kernel<<<256, 256, 0, st>>>(b.d_conv_weight, L.wp, L.cout, L.cin);
      if (L.wp2) pack_conv3x3_rowpair_kernel<<<64, 256, 0, st>>>(b.d_conv_weight, L.wp2, L.cout);
      CUDA_TRY(ctx, cudaMemcpyAsync(L.bias, b.d_conv_bias, L.cout * sizeof(float), cudaMemcpyDeviceToDevice, st));
      fold_bn_kernel<<<(L.cout + 127) / 128, 128, 0, st>>>(b.d_bn_gamma, b.d_bn_beta, b.d_bn_mean, b.d_bn_var,
                                                          w->bn_eps, L.scale, L.shift, L.cout);
    } else {
      if (!w->d_up_weight[L.up] || !w->d_up_bias[L.up]) return set_err(ctx, B200AL_ERR_ARG, "NULL upsample weight %d", L.up);
      pack_convT_kernel<<<256, 256, 0, st>>>(w->d_up_weight[L.up], L.wp, L.cin, L.cout);
      CUDA_TRY(ctx, cudaMemcpyAsync(L.bias, w->d_up_bias[L.up], L.cout * sizeof(float), cudaMemcpyDeviceToDevice, st));
      CUDA_TRY(ctx, cudaMemcpyAsync(L.scale, ctx->ones, L.cout * sizeof(float), cudaMemcpyDeviceToDevice, st));
      CUDA_TRY(ctx, cudaMemcpyAsync(L.shift, ctx->zeros, L.cout * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    L.p.scale = L.scale; L.p.bias = L.bias; L.p.shift = L.shift;
    ctx->launches += 2;
  }
  for (size_t i = 0; i < ctx->mix_layers.size(); ++i) {  // the scoring path's copies share the affine tables
    Layer& M = ctx->mix_layers[i];
    const Layer& L = ctx->layers[i];
    M.p.slope = L.p.slope; M.p.scale = L.scale; M.p.bias = L.bias; M.p.shift = L.shift;
    if (M.fmt < 2) continue;
    if (M.kind == KIND_CONV3) pack_conv3x3_f16_kernel<<<256, 256, 0, st>>>(w->blocks[M.block].d_conv_weight, M.wp16, M.cout, M.cin);
    else pack_convT_f16_kernel<<<256, 256, 0, st>>>(w->d_up_weight[M.up], M.wp16, M.cin, M.cout);
    ctx->launches++;
  }
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->head_w, w->d_head_weight, (size_t)ctx->K * 32 * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->head_b, w->d_head_bias, (size_t)ctx->K * sizeof(float), cudaMemcpyDeviceToDevice, st));
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaStreamSynchronize(st));  // the host copy of the first-conv constants must have landed
  ctx->weights_loaded = true;
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------
namespace {

// first conv over n images, each expanded to T samples (sample = image * T + t); mask nullable [n*T][2944]
// norm nullable [n][2] = per-image (mean, std) of the fused ZScoreNormalize
int run_first_conv(b200al_ctx* ctx, const void* d_images, int dtype, const float* norm, int n, int T, const float* mask,
                   cudaStream_t st) {
  // rows per block (sliding 3x3 window, one set-up per block): as many as leave >= 8 waves of 3 blocks per SM
  int rows = 8;
  while (rows > 1 && (long long)((ctx->H + rows - 1) / rows) * n < 8LL * 3 * ctx->sm_count) rows >>= 1;
  dim3 grid((ctx->W + 255) / 256, (ctx->H + rows - 1) / rows, n);
  const ActBuf& ob = ctx->bufs[E0A];
  const size_t smem = (size_t)T * 64 * sizeof(float);
  if (smem > 48 * 1024) return set_err(ctx, B200AL_ERR_ARG, "T = %d too large", T);
  const bool unit = ctx->slope >= 0.f && ctx->slope <= 1.f;
#define B200AL_FIRST(IMG, UNIT)                                                                                     \
  conv_first_kernel<IMG, UNIT><<<grid, 256, smem, st>>>((const IMG*)d_images, norm, T, ctx->H, ctx->W,              \
                                                        ctx->first_consts, mask, B200AL_MASK_CHANNELS, ctx->slope, \
                                                        ob.ptr, 2 * ob.c_total, 0, ob.c_total, ctx->nsplit == 3, rows)
  if (dtype == B200AL_U8) {
    if (unit) B200AL_FIRST(uint8_t, true); else B200AL_FIRST(uint8_t, false);
  } else if (dtype == B200AL_F32) {
    if (unit) B200AL_FIRST(float, true); else B200AL_FIRST(float, false);
  } else {
    return set_err(ctx, B200AL_ERR_ARG, "bad image dtype %d", dtype);
  }
#undef B200AL_FIRST
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return 0;
}

// layers [0, n_layers) of ctx->layers on S samples; mask nullable [S][2944]
// head_logits non-null: the last layer applies the fused seg_output head and writes logits there
// mixed: the scoring path's layer list (fp16 group in the deep layers) when the context was created with one
int run_layers(b200al_ctx* ctx, int n_layers, int S, const float* mask, cudaStream_t st, float* head_logits = nullptr,
               bool mixed = false) {
  std::vector<Layer>& layers = (mixed && !ctx->mix_layers.empty()) ? ctx->mix_layers : ctx->layers;
  for (int i = 0; i < n_layers; ++i) {
    Layer& L = layers[i];
    ConvTcParams p = L.p;
    p.mask = (mask && L.kind == KIND_CONV3) ? mask + L.mask_off : nullptr;
    int rc = launch_conv(ctx, L, L.tmA, L.tmO, p, S, st, i == n_layers - 1 ? head_logits : nullptr);
    if (rc) return rc;
  }
  return 0;
}

int prepare_masks(b200al_ctx* ctx, int n, int T, float drop_p, uint64_t seed, int64_t base_pos,
                  const float* d_user_masks, const float** mask_out, cudaStream_t st) {
  *mask_out = nullptr;
  if (d_user_masks) { *mask_out = d_user_masks; return 0; }
  if (drop_p <= 0.f) return 0;
  if (drop_p >= 1.f) return set_err(ctx, B200AL_ERR_ARG, "drop_p must be < 1");
  const uint32_t thr = (uint32_t)floor((double)drop_p * 16777216.0);
  const float keep_scale = 1.0f / (1.0f - drop_p);
  const long long total = (long long)n * T * B200AL_MASK_CHANNELS;
  const int blocks = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  mc_mask_kernel<<<blocks, 256, 0, st>>>(ctx->mask, n * T, T, B200AL_MASK_CHANNELS, seed, (long long)base_pos, thr,
                                        keep_scale);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  *mask_out = ctx->mask;
  return 0;
}

template <int K>
void launch_head(b200al_ctx* ctx, int n, int T, int mode, float smooth, float* logits, cudaStream_t st) {
  const ActBuf& fb = ctx->bufs[D3B];
  dim3 grid(ctx->partial_blocks, n);
  head_uncertainty_kernel<K><<<grid, 256, 0, st>>>(fb.ptr, 2 * fb.c_total, 0, fb.c_total, ctx->nsplit == 3,
                                                  ctx->H * ctx->W, T, ctx->head_w, ctx->head_b, mode, smooth, logits,
                                                  ctx->partial);
}

template <int K>
void launch_unc(const float* logits, int n, int T, int HW, int mode, float smooth, double* partial, int nblk,
                cudaStream_t st) {
  dim3 grid(nblk, n);
  uncertainty_logits_kernel<K><<<grid, 256, 0, st>>>(logits, T, HW, mode, smooth, partial);
}

// dice-uncertainty over fp32 logits [n][T][K][HW]: 3K sums per image, then the DiceLoss formula (uncertainty.cuh)
template <int K>
void launch_dice(const float* logits, int n, int T, int HW, float smooth, double* partial, float* scores,
                 cudaStream_t st) {
  int nblk = (HW / 4 + 255) / 256;
  if (nblk > kDiceBlocks) nblk = kDiceBlocks;
  dice_logits_kernel<K><<<dim3(nblk, n), 256, 0, st>>>(logits, T, HW, partial);
  dice_finalize_kernel<K><<<(n + 127) / 128, 128, 0, st>>>(partial, nblk, smooth, scores, n);
}

#define DISPATCH_K(K_, CALL)                         \
  switch (K_) {                                      \
    case 2: CALL(2); break;                          \
    case 3: CALL(3); break;                          \
    case 4: CALL(4); break;                          \
    case 5: CALL(5); break;                          \
    case 6: CALL(6); break;                          \
    case 7: CALL(7); break;                          \
    case 8: CALL(8); break;                          \
    default: break;                                  \
  }

}  // namespace

extern "C" int b200al_prep_images(b200al_ctx* ctx, const uint8_t* d_raw, const int64_t* d_offsets, const int32_t* d_hw,
                                  int n, int max_h, int max_w, float* d_images_out, float* d_norm_out, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n == 0) return B200AL_OK;
  if (!d_raw || !d_offsets || !d_hw || !d_images_out || n < 0 || max_h < 1 || max_w < 1)
    return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (ctx->W > 1024) return set_err(ctx, B200AL_ERR_ARG, "prep supports W <= 1024");
  const float sc_h = (float)max_h / (float)ctx->H, sc_w = (float)max_w / (float)ctx->W;
  const float sc = sc_h > sc_w ? sc_h : sc_w;
  const int taps = 2 * (int)ceilf(sc > 1.f ? sc : 1.f) + 1;
  if (taps > kPrepMaxTaps)
    return set_err(ctx, B200AL_ERR_ARG, "source image %d x %d is more than 17x the model input", max_h, max_w);
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int bands = (ctx->H + kPrepBand - 1) / kPrepBand;
  const size_t smem = ((size_t)taps * ctx->W + (size_t)kPrepBand * taps) * sizeof(float);
  if (smem > 200 * 1024) return set_err(ctx, B200AL_ERR_ARG, "prep filter table does not fit shared memory");
  static_assert(kPrepBand <= 32, "band rows are set up by the first warp");
  CUDA_TRY(ctx, cudaFuncSetAttribute(prep_resize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int i0 = 0; i0 < n; i0 += ctx->max_samples) {  // the partials buffer holds max_samples images
    const int nn = (n - i0) < ctx->max_samples ? (n - i0) : ctx->max_samples;
    double* part = d_norm_out ? ctx->prep_partials : nullptr;
    prep_resize_kernel<<<dim3(bands, nn), ctx->W, smem, st>>>(d_raw, d_offsets + i0, d_hw + 2 * (size_t)i0, ctx->H,
                                                              ctx->W, taps, d_images_out + (size_t)i0 * ctx->H * ctx->W,
                                                              part);
    ctx->launches++;
    if (d_norm_out) {
      prep_stats_kernel<<<(nn + 127) / 128, 128, 0, st>>>(part, bands, ctx->H * ctx->W, d_norm_out + 2 * (size_t)i0, nn);
      ctx->launches++;
    }
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_unet_scores(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n,
                                  int mode, int T,
                                  float drop_p, uint64_t mc_seed, int64_t base_pos, const float* d_drop_masks,
                                  float smooth, float* d_scores_out, float* d_logits_out, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (!ctx->weights_loaded) return set_err(ctx, B200AL_ERR_STATE, "b200al_load_weights has not been called");
  if (n == 0) return B200AL_OK;
  if (!d_images || !d_scores_out || n < 0 || T < 1) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if ((long long)n * T > ctx->max_samples)
    return set_err(ctx, B200AL_ERR_ARG, "n*T = %lld exceeds max_samples = %d", (long long)n * T, ctx->max_samples);
  if (mode < 0 || mode > 3) return set_err(ctx, B200AL_ERR_ARG, "bad score mode %d", mode);
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (mode == B200AL_DICE && !ctx->fuse_head && !d_logits_out && !ctx->logits_buf) {
    // unfused head (K > 4): the dice reduction reads logits, so they need a home; allocated on first use
    if (int rc0 = dev_alloc(ctx, &ctx->logits_buf, (size_t)ctx->max_samples * ctx->K * ctx->H * ctx->W)) return rc0;
  }
  const float* mask = nullptr;
  int rc = prepare_masks(ctx, n, T, drop_p, mc_seed, base_pos, d_drop_masks, &mask, st);
  if (rc) return rc;
  const int S = n * T;
  if ((rc = run_first_conv(ctx, d_images, image_dtype, d_norm, n, T, mask, st))) return rc;
  int nblk = ctx->partial_blocks;
  if (ctx->fuse_head) {
    // last decoder block + seg_output fused: logits [S][K][HW] fp32, then the HBM-bound softmax/score reduction
    float* lg = d_logits_out ? d_logits_out : ctx->logits_buf;
    if ((rc = run_layers(ctx, (int)ctx->layers.size(), S, mask, st, lg, true))) return rc;
    const int HW = ctx->H * ctx->W;
    if ((HW / 4 + 255) / 256 < nblk) nblk = (HW / 4 + 255) / 256;
    if (mode == B200AL_DICE) {
#define CALL_DICE(KK) launch_dice<KK>(lg, n, T, HW, smooth, ctx->dice_partial, d_scores_out, st)
      DISPATCH_K(ctx->K, CALL_DICE)
#undef CALL_DICE
      ctx->launches += 2;
      CUDA_TRY(ctx, cudaGetLastError());
      return B200AL_OK;
    }
#define CALL_UNC(KK) launch_unc<KK>(lg, n, T, HW, mode, smooth, ctx->partial, nblk, st)
    DISPATCH_K(ctx->K, CALL_UNC)
#undef CALL_UNC
  } else {
    if ((rc = run_layers(ctx, (int)ctx->layers.size(), S, mask, st, nullptr, true))) return rc;
    float* lg = mode == B200AL_DICE ? (d_logits_out ? d_logits_out : ctx->logits_buf) : d_logits_out;
#define CALL_HEAD(KK) launch_head<KK>(ctx, n, T, mode == B200AL_DICE ? 0 : mode, smooth, lg, st)
    DISPATCH_K(ctx->K, CALL_HEAD)
#undef CALL_HEAD
    if (mode == B200AL_DICE) {
#define CALL_DICE(KK) launch_dice<KK>(lg, n, T, ctx->H * ctx->W, smooth, ctx->dice_partial, d_scores_out, st)
      DISPATCH_K(ctx->K, CALL_DICE)
#undef CALL_DICE
      ctx->launches += 3;
      CUDA_TRY(ctx, cudaGetLastError());
      return B200AL_OK;
    }
  }
  uncertainty_finalize_kernel<<<(n + 127) / 128, 128, 0, st>>>(ctx->partial, nblk, ctx->H * ctx->W, d_scores_out, n);
  ctx->launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_unet_embed(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n,
                                 float* d_emb_out, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (!ctx->weights_loaded) return set_err(ctx, B200AL_ERR_STATE, "b200al_load_weights has not been called");
  if (n == 0) return B200AL_OK;
  if (!d_images || !d_emb_out || n < 0) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (n > ctx->max_samples) return set_err(ctx, B200AL_ERR_ARG, "n = %d exceeds max_samples = %d", n, ctx->max_samples);
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int rc;
  if ((rc = run_first_conv(ctx, d_images, image_dtype, d_norm, n, 1, nullptr, st))) return rc;
  if ((rc = run_layers(ctx, kEncoderLayers, n, nullptr, st))) return rc;
  const ActBuf& bb = ctx->bufs[E4B];
  avgpool_kernel<<<n, 512, 0, st>>>(bb.ptr, (ctx->H >> 4) * (ctx->W >> 4), 2 * bb.c_total, 0, bb.c_total,
                                    ctx->nsplit == 3, bb.c_total, d_emb_out);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

namespace {
template <int K>
void launch_badge(b200al_ctx* ctx, int n, float smooth, int do_bg, float* out, int out_stride, cudaStream_t st) {
  const ActBuf& fb = ctx->bufs[D3B];
  const int HW = ctx->H * ctx->W;
  dim3 grid(kBadgeBlocks, n);
  badge_stats_kernel<K><<<grid, 256, 0, st>>>(fb.ptr, 2 * fb.c_total, 0, fb.c_total, ctx->nsplit == 3, HW, ctx->head_w,
                                             ctx->head_b, ctx->badge_stats);
  badge_coef_kernel<K><<<(n + 127) / 128, 128, 0, st>>>(ctx->badge_stats, kBadgeBlocks, smooth, do_bg, ctx->badge_coef, n);
  badge_grad_kernel<K><<<grid, 256, 0, st>>>(fb.ptr, 2 * fb.c_total, 0, fb.c_total, ctx->nsplit == 3, HW, ctx->head_w,
                                            ctx->head_b, ctx->badge_coef, ctx->badge_grad);
  badge_reduce_kernel<<<(n * out_stride + 255) / 256, 256, 0, st>>>(ctx->badge_grad, kBadgeBlocks, K * 32, out_stride,
                                                                   out, n);
}
}  // namespace

extern "C" int b200al_badge_embed(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n,
                                  float dice_smooth, int dice_do_bg, float* d_emb_out, int out_stride, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (!ctx->weights_loaded) return set_err(ctx, B200AL_ERR_STATE, "b200al_load_weights has not been called");
  if (n == 0) return B200AL_OK;
  if (!d_images || !d_emb_out || n < 0) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (n > ctx->max_samples) return set_err(ctx, B200AL_ERR_ARG, "n = %d exceeds max_samples = %d", n, ctx->max_samples);
  if (ctx->K > kBadgeMaxK) return set_err(ctx, B200AL_ERR_UNSUPPORTED, "BADGE embedding supports up to %d classes", kBadgeMaxK);
  if (out_stride < ctx->K * 32) return set_err(ctx, B200AL_ERR_ARG, "out_stride must be >= K*32 = %d", ctx->K * 32);
  if (!dice_do_bg && ctx->K < 2) return set_err(ctx, B200AL_ERR_ARG, "do_bg = 0 needs K >= 2");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int rc;
  if (!ctx->badge_stats) {  // scratch of the BADGE path, allocated on first use, freed with the context
    if ((rc = dev_alloc(ctx, &ctx->badge_stats, (size_t)ctx->max_samples * kBadgeBlocks * 3 * kBadgeMaxK))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->badge_grad, (size_t)ctx->max_samples * kBadgeBlocks * 32 * kBadgeMaxK))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->badge_coef, (size_t)ctx->max_samples * 2 * kBadgeMaxK))) return rc;
  }
  if ((rc = run_first_conv(ctx, d_images, image_dtype, d_norm, n, 1, nullptr, st))) return rc;
  // all 21 conv layers WITHOUT the fused head: the last block leaves its 32-channel feature map in HBM
  if ((rc = run_layers(ctx, (int)ctx->layers.size(), n, nullptr, st))) return rc;
  switch (ctx->K) {
    case 2: launch_badge<2>(ctx, n, dice_smooth, dice_do_bg, d_emb_out, out_stride, st); break;
    case 3: launch_badge<3>(ctx, n, dice_smooth, dice_do_bg, d_emb_out, out_stride, st); break;
    case 4: launch_badge<4>(ctx, n, dice_smooth, dice_do_bg, d_emb_out, out_stride, st); break;
    default: break;
  }
  ctx->launches += 4;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_uncertainty_from_logits(b200al_ctx* ctx, const float* d_logits, int n, int T, int K, int HW,
                                              int mode, float smooth, float* d_scores_out, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n == 0) return B200AL_OK;
  if (!d_logits || !d_scores_out || n < 0 || T < 1 || K < 2 || K > kMaxClasses || HW < 4 || HW % 4)
    return set_err(ctx, B200AL_ERR_ARG, "bad argument (need 2 <= K <= %d, HW %% 4 == 0)", kMaxClasses);
  if (mode < 0 || mode > 3) return set_err(ctx, B200AL_ERR_ARG, "bad score mode %d", mode);
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int nblk = (HW / 4 + 255) / 256;
  if (nblk > ctx->partial_blocks) nblk = ctx->partial_blocks;
  for (int i0 = 0; i0 < n; i0 += ctx->max_samples) {
    const int nn = (n - i0) < ctx->max_samples ? (n - i0) : ctx->max_samples;
    const float* lg = d_logits + (size_t)i0 * T * K * HW;
    if (mode == B200AL_DICE) {
#define CALL_DICE(KK) launch_dice<KK>(lg, nn, T, HW, smooth, ctx->dice_partial, d_scores_out + i0, st)
      DISPATCH_K(K, CALL_DICE)
#undef CALL_DICE
      ctx->launches += 2;
      continue;
    }
#define CALL_UNC(KK) launch_unc<KK>(lg, nn, T, HW, mode, smooth, ctx->partial, nblk, st)
    DISPATCH_K(K, CALL_UNC)
#undef CALL_UNC
    uncertainty_finalize_kernel<<<(nn + 127) / 128, 128, 0, st>>>(ctx->partial, nblk, HW, d_scores_out + i0, nn);
    ctx->launches += 2;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// workspace (scratch of b200al_topk / b200al_kcenter_fused)
// ---------------------------------------------------------------------------
extern "C" size_t b200al_workspace_bytes_topk(int n) {
  return n > 0 ? (size_t)n * 2 * sizeof(unsigned long long) : 0;
}
extern "C" size_t b200al_workspace_bytes_kcenter_fused(int n, int D1, int D2, int metric) {
  if (n < 1 || D1 < 1 || D2 < 1) return 0;
  const size_t nwarps = (size_t)148 * 8 * 8 * 2;  // upper bound of the partial-sum slots on any sm_100 part
  const size_t norm = metric == B200AL_COSINE ? (size_t)n * ((size_t)D1 + (size_t)D2) * 4 : 0;
  return (size_t)n * 8 + nwarps * 8 + 16 + (((size_t)n + 15) & ~(size_t)15) + norm + 256;
}
// Grows the context's workspace to at least `bytes` (synchronises the device when it has to reallocate; a no-op
// otherwise).  The one place where this scratch is allocated: call it at set-up, not between hot calls.
extern "C" int b200al_workspace_reserve(b200al_ctx* ctx, size_t bytes) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (bytes <= ctx->ws_bytes) return B200AL_OK;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  if (ctx->ws) cudaFree(ctx->ws);
  ctx->ws = nullptr;
  ctx->ws_bytes = 0;
  CUDA_TRY(ctx, cudaMalloc(&ctx->ws, bytes));
  ctx->ws_bytes = bytes;
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// rank & cut
// ---------------------------------------------------------------------------
extern "C" int b200al_topk(b200al_ctx* ctx, const float* d_scores, int n, int k, int64_t base_pos,
                           float* d_out_scores, int64_t* d_out_pos, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n < 0 || k < 0) return set_err(ctx, B200AL_ERR_ARG, "negative size");
  if (k > n) k = n;
  if (k == 0) return B200AL_OK;
  if (!d_scores || !d_out_scores || !d_out_pos) return set_err(ctx, B200AL_ERR_ARG, "NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t need = b200al_workspace_bytes_topk(n);
  if (need > ctx->ws_bytes)
    return set_err(ctx, B200AL_ERR_STATE, "b200al_topk: workspace of %zu bytes needed, %zu reserved - call "
                   "b200al_workspace_reserve(ctx, b200al_workspace_bytes_topk(n)) first", need, ctx->ws_bytes);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(ctx->ws);
  unsigned long long* cand = keys + n;
  const int blocks = (n + 255) / 256 < ctx->sm_count * 8 ? (n + 255) / 256 : ctx->sm_count * 8;
  topk_keys_kernel<<<blocks, 256, 0, st>>>(d_scores, n, keys, ctx->tk_state, (unsigned)k);
  for (int pass = 0; pass < 8; ++pass) {
    topk_hist_kernel<<<blocks, 256, 0, st>>>(keys, n, ctx->tk_state, pass);
    topk_scan_kernel<<<1, 32, 0, st>>>(ctx->tk_state, pass);
  }
  topk_compact_kernel<<<blocks, 256, 0, st>>>(keys, n, ctx->tk_state, cand, (unsigned)k);
  topk_rank_kernel<<<(k + 255) / 256, 256, 0, st>>>(cand, (unsigned)k, (long long)base_pos, d_out_scores,
                                                   (long long*)d_out_pos);
  ctx->launches += 19;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// rank & cut without allocation (include/b200al.h: b200al_topk_static)
// ---------------------------------------------------------------------------
namespace {
// layout: keys [MAX_N] u64 | candidates [MAX_N] u64 | TopkState
constexpr size_t kTopkStaticKeysBytes = (size_t)B200AL_TOPK_STATIC_MAX_N * sizeof(unsigned long long);
constexpr size_t kTopkStaticBytes = 2 * kTopkStaticKeysBytes + ((sizeof(TopkState) + 255) / 256) * 256;
}  // namespace
// File scope, not in a namespace: in HOST code this name is only a stand-in for the symbol (its host address is not a
// device address); the device address is resolved per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_topk_static_scratch[kTopkStaticBytes];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {nullptr};  // device address of g_topk_static_scratch on device i
int g_static_sms[kMaxDevices] = {0};

// once per device, with that device current
int topk_static_setup(int device, void** scratch, int* sms) {
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_topk_static_scratch);
    if (e != cudaSuccess || !ptr)
      return set_err(nullptr, B200AL_ERR_CUDA, "cudaGetSymbolAddress(topk static scratch) on device %d: %s", device,
                     cudaGetErrorString(e));
    int n_sm = 0;
    e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || n_sm < 1)
      return set_err(nullptr, B200AL_ERR_CUDA, "cudaDeviceGetAttribute(SM count) on device %d: %s", device,
                     cudaGetErrorString(e));
    g_static_sms[device] = n_sm;
    g_static_ptr[device] = ptr;
  }
  *scratch = g_static_ptr[device];
  *sms = g_static_sms[device];
  return 0;
}
}  // namespace

extern "C" int b200al_topk_static(int device, const float* d_scores, int n, int k, int64_t base_pos,
                                  float* d_out_scores, int64_t* d_out_pos, void* stream) {
  if (n < 0 || k < 0) return set_err(nullptr, B200AL_ERR_ARG, "negative size");
  if (n > B200AL_TOPK_STATIC_MAX_N)
    return set_err(nullptr, B200AL_ERR_ARG,
                   "b200al_topk_static: n = %d exceeds the static scratch (B200AL_TOPK_STATIC_MAX_N = %d); use "
                   "b200al_topk", n, B200AL_TOPK_STATIC_MAX_N);
  if (device < 0 || device >= kMaxDevices) return set_err(nullptr, B200AL_ERR_ARG, "device %d out of range", device);
  if (k > n) k = n;
  if (k == 0) return B200AL_OK;
  if (!d_scores || !d_out_scores || !d_out_pos) return set_err(nullptr, B200AL_ERR_ARG, "NULL pointer");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(nullptr, cudaSetDevice(device));
  void* scratch = nullptr;
  int sms = 0;
  if (int rc = topk_static_setup(device, &scratch, &sms)) return rc;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(scratch);
  unsigned long long* cand = keys + B200AL_TOPK_STATIC_MAX_N;
  TopkState* state = reinterpret_cast<TopkState*>(reinterpret_cast<unsigned char*>(scratch) + 2 * kTopkStaticKeysBytes);
  // same launches as b200al_topk; topk_keys_kernel resets the state (prefix, remaining, count, histogram) and every
  // key / candidate slot that is read is written first, so the previous call's contents never matter
  const int blocks = (n + 255) / 256 < sms * 8 ? (n + 255) / 256 : sms * 8;
  topk_keys_kernel<<<blocks, 256, 0, st>>>(d_scores, n, keys, state, (unsigned)k);
  for (int pass = 0; pass < 8; ++pass) {
    topk_hist_kernel<<<blocks, 256, 0, st>>>(keys, n, state, pass);
    topk_scan_kernel<<<1, 32, 0, st>>>(state, pass);
  }
  topk_compact_kernel<<<blocks, 256, 0, st>>>(keys, n, state, cand, (unsigned)k);
  topk_rank_kernel<<<(k + 255) / 256, 256, 0, st>>>(cand, (unsigned)k, (long long)base_pos, d_out_scores,
                                                   (long long*)d_out_pos);
  CUDA_TRY(nullptr, cudaGetLastError());
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// k-center greedy
// ---------------------------------------------------------------------------
namespace {
template <int NV>
void launch_kc_init(const float* emb, int n, const float* centres, int L, int metric, int criteria, void* state,
                    int accumulate, int blocks, cudaStream_t st) {
  if (criteria == B200AL_MEAN)
    kcenter_init_mean_kernel<NV, (NV <= 4 ? 4 : NV <= 8 ? 2 : 1)><<<blocks, 256, 0, st>>>(emb, n, centres, L, metric, (double*)state, accumulate);
  else
    kcenter_init_kernel<NV, (NV <= 4 ? 4 : NV <= 8 ? 2 : 1)><<<blocks, 256, 0, st>>>(emb, n, centres, L, metric, (float*)state, accumulate);
}
template <int NV>
void launch_kc_step(const float* emb, int n, const float* centre, int do_update, int metric, int criteria, void* state,
                    double inv_count, const uint8_t* taken, float* pv, int* pi, int blocks, cudaStream_t st) {
  if (criteria == B200AL_MEAN)
    kcenter_step_mean_kernel<NV><<<blocks, 256, 0, st>>>(emb, n, centre, do_update, metric, (double*)state, inv_count, taken, pv, pi);
  else
    kcenter_step_kernel<NV><<<blocks, 256, 0, st>>>(emb, n, centre, do_update, metric, (float*)state, taken, pv, pi);
}
int check_dim(b200al_ctx* ctx, int D) {
  if (D != 128 && D != 256 && D != 512 && D != 1024 && D != 1536 && D != 2048)
    return set_err(ctx, B200AL_ERR_UNSUPPORTED, "row width %d not in {128,256,512,1024,1536,2048}", D);
  return 0;
}
}  // namespace

extern "C" int b200al_normalize_rows(b200al_ctx* ctx, const float* d_in, float* d_out, int n, int D, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n == 0) return B200AL_OK;
  if (!d_in || !d_out || n < 0 || D < 1) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  normalize_rows_kernel<<<(n + 7) / 8, 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n, D);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_kcenter_init(b200al_ctx* ctx, const float* d_emb, int n, int D, const float* d_centres, int L,
                                   int metric, int criteria, void* d_state, int accumulate, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n == 0) return B200AL_OK;
  if (!d_emb || !d_centres || !d_state || n < 0) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (L < 1) return set_err(ctx, B200AL_ERR_ARG, "k-center needs at least one initial centre (L >= 1)");
  if (metric != B200AL_L2 && metric != B200AL_COSINE) return set_err(ctx, B200AL_ERR_ARG, "bad metric");
  if (criteria != B200AL_MIN && criteria != B200AL_MEAN) return set_err(ctx, B200AL_ERR_ARG, "bad criteria");
  if (int rc = check_dim(ctx, D)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int blocks = (n + 31) / 32;  // 8 warps x 4 rows
  if (blocks > ctx->sm_count * 8) blocks = ctx->sm_count * 8;
  switch (D / 128) {
    case 1: launch_kc_init<1>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
    case 2: launch_kc_init<2>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
    case 4: launch_kc_init<4>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
    case 8: launch_kc_init<8>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
    case 12: launch_kc_init<12>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
    case 16: launch_kc_init<16>(d_emb, n, d_centres, L, metric, criteria, d_state, accumulate, blocks, st); break;
  }
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_kcenter_fused(b200al_ctx* ctx, const float* d_emb, int D1, const float* d_loaded, int D2, int n,
                                    int n_init, int metric, int criteria, float loaded_weight, float emb_weight,
                                    int budget, int64_t* d_picks, float* d_pick_vals, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (budget == 0) return B200AL_OK;
  if (!d_emb || !d_loaded || !d_picks || n < 1 || budget < 0 || D1 < 1 || D2 < 1 || D1 > 1024)
    return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (n_init < 1 || n_init > n) return set_err(ctx, B200AL_ERR_ARG, "k-center needs 1 <= n_init <= n initial centres");
  if (metric != B200AL_L2 && metric != B200AL_COSINE) return set_err(ctx, B200AL_ERR_ARG, "bad metric");
  if (criteria != B200AL_MIN && criteria != B200AL_MEAN) return set_err(ctx, B200AL_ERR_ARG, "bad criteria");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int sblocks = (n + 7) / 8;
  if (sblocks > ctx->sm_count * 8) sblocks = ctx->sm_count * 8;
  const int nwarps = sblocks * 8;
  // scratch: state [n] doubles | partial [nwarps] doubles | sums [2] floats | taken [n] | normalised copies (cosine)
  const size_t norm_elems = metric == B200AL_COSINE ? (size_t)n * (D1 + D2) : 0;
  const size_t bytes = (size_t)n * 8 + (size_t)nwarps * 8 + 16 + (((size_t)n + 15) & ~(size_t)15) + norm_elems * 4;
  if (bytes > b200al_workspace_bytes_kcenter_fused(n, D1, D2, metric))
    return set_err(ctx, B200AL_ERR_CUDA, "kcenter_fused: workspace size formula out of date");
  if (bytes > ctx->ws_bytes)
    return set_err(ctx, B200AL_ERR_STATE, "b200al_kcenter_fused: workspace of %zu bytes needed, %zu reserved - call "
                   "b200al_workspace_reserve(ctx, b200al_workspace_bytes_kcenter_fused(n, D1, D2, metric)) first",
                   bytes, ctx->ws_bytes);
  char* scratch = ctx->ws;
  void* state = scratch;
  double* partial = (double*)(scratch + (size_t)n * 8);
  float* sums = (float*)(partial + nwarps);
  uint8_t* taken = (uint8_t*)(sums + 4);
  float* n1 = (float*)(taken + (((size_t)n + 15) & ~(size_t)15));
  float* n2 = n1 + (size_t)n * D1;
  FusedFeats f;
  f.e1 = d_emb; f.e2 = d_loaded; f.d1 = D1; f.d2 = D2; f.metric = metric;
  f.s1 = f.s2 = 1.f;
  f.w2 = loaded_weight;
  f.w1 = emb_weight;
  if (metric == B200AL_COSINE) {
    normalize_rows_kernel<<<(n + 7) / 8, 256, 0, st>>>(d_emb, n1, n, D1);
    normalize_rows_kernel<<<(n + 7) / 8, 256, 0, st>>>(d_loaded, n2, n, D2);
    f.e1 = n1; f.e2 = n2;
    ctx->launches += 2;
  }
  // the two global sums (coreset_selector.py:145,153): sums[0] = model embeddings, sums[1] = loaded features
  pair_distance_sum_kernel<<<sblocks, 256, 0, st>>>(f.e1, n, D1, metric, partial);
  sum_partials_kernel<<<1, 32, 0, st>>>(partial, nwarps, sums);
  pair_distance_sum_kernel<<<sblocks, 256, 0, st>>>(f.e2, n, D2, metric, partial);
  sum_partials_kernel<<<1, 32, 0, st>>>(partial, nwarps, sums + 1);
  cudaMemsetAsync(taken, 0, n, st);
  cudaMemsetAsync(taken, 1, n_init, st);
  kcenter_fused_init_kernel<<<sblocks, 256, 0, st>>>(f, sums, n, n_init, criteria, state);
  ctx->launches += 5;
  int blocks = (n + 63) / 64;
  if (blocks > ctx->kc_blocks) blocks = ctx->kc_blocks;
  if (blocks < 1) blocks = 1;
  for (int step = 0; step < budget; ++step) {
    const double inv = 1.0 / (double)(n_init + step);
    kcenter_fused_step_kernel<<<blocks, 256, 0, st>>>(f, sums, n, (const long long*)d_picks, step, criteria, state, inv,
                                                     taken, ctx->kc_part_val, ctx->kc_part_idx);
    kcenter_pick_kernel<<<1, 256, 0, st>>>(ctx->kc_part_val, ctx->kc_part_idx, blocks, f.e1, D1, taken, ctx->kc_centre,
                                          (long long*)d_picks, d_pick_vals, step);
    ctx->launches += 2;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_standardize_rows(b200al_ctx* ctx, const float* d_in, float* d_out, int n, int D, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n == 0) return B200AL_OK;
  if (!d_in || !d_out || n < 0 || D < 2) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  standardize_rows_kernel<<<(n + 7) / 8, 256, 0, st>>>(d_in, d_out, n, D);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" size_t b200al_kmeans_pp_scratch_bytes(int n, int n_local_trials) {
  if (n < 1 || n_local_trials < 1) return 0;
  const size_t blocks = 148 * 8;
  return ((size_t)n * (n_local_trials + 1)) * sizeof(float) + blocks * n_local_trials * sizeof(double) + 256 +
         (size_t)n * sizeof(double) + 16;
}

namespace {
template <int NV>
void launch_kmpp_dist(const float* X, const double* xx, const float* w, int n, const int* cand, int trials,
                      const float* closest, int first, float* cand_d, double* part, int blocks, cudaStream_t st) {
  const size_t smem = (size_t)trials * NV * 128 * sizeof(double);
  cudaFuncSetAttribute(kmpp_dist_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kmpp_dist_kernel<NV><<<blocks, 256, smem, st>>>(X, xx, w, n, cand, trials, closest, first, cand_d, part);
}
}  // namespace

extern "C" int b200al_kmeans_pp(b200al_ctx* ctx, const float* d_X, int n, int D, const float* d_weight, int first_centre,
                                const double* d_uniform, int n_clusters, int n_local_trials, void* d_scratch,
                                int64_t* d_picks, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (n_clusters == 0) return B200AL_OK;
  if (!d_X || !d_weight || !d_scratch || !d_picks || n < 1 || n_clusters < 0 || first_centre < 0 || first_centre >= n)
    return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (n_clusters > n) return set_err(ctx, B200AL_ERR_ARG, "n_samples=%d should be >= n_clusters=%d", n, n_clusters);
  if (n_local_trials < 1 || n_local_trials > kKmppMaxTrials)
    return set_err(ctx, B200AL_ERR_ARG, "n_local_trials must be in [1, %d]", kKmppMaxTrials);
  if (n_clusters > 1 && !d_uniform) return set_err(ctx, B200AL_ERR_ARG, "d_uniform is NULL");
  if (int rc = check_dim(ctx, D)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  int blocks = (n + 63) / 64;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const int trials = n_local_trials;
  // scratch: closest [n] | cand_d [trials][n] | part [trials][blocks] doubles | pot, best, cand[16]
  float* closest = (float*)d_scratch;
  float* cand_d = closest + n;
  size_t off = ((size_t)n * (trials + 1) * sizeof(float) + 15) & ~(size_t)15;
  double* part = (double*)((char*)d_scratch + off);
  off += (size_t)148 * 8 * trials * sizeof(double);
  float* pot = (float*)((char*)d_scratch + off);
  int* best = (int*)(pot + 1);
  int* cand = best + 1;
  off = (off + 256 + 15) & ~(size_t)15;
  double* xx = (double*)((char*)d_scratch + off);  // [n] squared row norms
  long long* picks = (long long*)d_picks;
  if ((size_t)trials * D * sizeof(double) > 200 * 1024)
    return set_err(ctx, B200AL_ERR_ARG, "n_local_trials * D too large for the candidate tile");
  switch (D / 128) {
    case 1: kmpp_norms_kernel<1><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
    case 2: kmpp_norms_kernel<2><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
    case 4: kmpp_norms_kernel<4><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
    case 8: kmpp_norms_kernel<8><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
    case 12: kmpp_norms_kernel<12><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
    case 16: kmpp_norms_kernel<16><<<blocks, 256, 0, st>>>(d_X, n, xx); break;
  }
  ctx->launches++;
#define KMPP_DIST(TR, FIRST)                                                                                   \
  switch (D / 128) {                                                                                           \
    case 1: launch_kmpp_dist<1>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
    case 2: launch_kmpp_dist<2>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
    case 4: launch_kmpp_dist<4>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
    case 8: launch_kmpp_dist<8>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
    case 12: launch_kmpp_dist<12>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
    case 16: launch_kmpp_dist<16>(d_X, xx, d_weight, n, cand, TR, closest, FIRST, cand_d, part, blocks, st); break;   \
  }
  kmpp_set_first_kernel<<<1, 1, 0, st>>>(cand, first_centre);
  KMPP_DIST(1, 1)
  kmpp_pick_kernel<<<1, 32, 0, st>>>(part, blocks, cand, 1, pot, best, picks);
  kmpp_commit_kernel<<<blocks, 256, 0, st>>>(cand_d, best, n, closest);
  ctx->launches += 4;
  for (int c = 1; c < n_clusters; ++c) {
    kmpp_sample_kernel<<<1, 1024, 0, st>>>(d_weight, closest, n, d_uniform + (size_t)(c - 1) * trials, trials, pot, cand);
    KMPP_DIST(trials, 0)
    kmpp_pick_kernel<<<1, 32 * trials, 0, st>>>(part, blocks, cand, trials, pot, best, picks + c);
    kmpp_commit_kernel<<<blocks, 256, 0, st>>>(cand_d, best, n, closest);
    ctx->launches += 4;
  }
#undef KMPP_DIST
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

namespace {
template <int NV>
int launch_kc_persist(b200al_ctx* ctx, KcPersistParams p, cudaStream_t st) {
  int occ = 0;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kcenter_persist_kernel<NV>, 256, 0));
  if (occ < 1) return set_err(ctx, B200AL_ERR_CUDA, "persistent k-center kernel does not fit an SM");
  int blocks = (p.n_local + 63) / 64;  // >= 8 rows per warp
  const int resident = ctx->sm_count * (occ > 4 ? 4 : occ);  // a cooperative grid must be co-resident
  if (blocks > resident) blocks = resident;
  if (blocks > ctx->kc_blocks) blocks = ctx->kc_blocks;
  if (blocks < 1) blocks = 1;
  p.part_val = ctx->kc_pp_val;
  p.part_idx = ctx->kc_pp_idx;
  void* args[] = {&p};
  CUDA_TRY(ctx, cudaLaunchCooperativeKernel((const void*)kcenter_persist_kernel<NV>, dim3(blocks), dim3(256), args, 0, st));
  ctx->launches++;
  return 0;
}
int run_kc_persist(b200al_ctx* ctx, int D, const KcPersistParams& p, cudaStream_t st) {
  switch (D / 128) {
    case 1: return launch_kc_persist<1>(ctx, p, st);
    case 2: return launch_kc_persist<2>(ctx, p, st);
    case 4: return launch_kc_persist<4>(ctx, p, st);
    case 8: return launch_kc_persist<8>(ctx, p, st);
    case 12: return launch_kc_persist<12>(ctx, p, st);
    case 16: return launch_kc_persist<16>(ctx, p, st);
  }
  return set_err(ctx, B200AL_ERR_UNSUPPORTED, "row width %d", D);
}
}  // namespace

extern "C" size_t b200al_kcenter_mailbox_bytes(int world) {
  return world > 0 && world <= kKcMaxWorld ? (size_t)4 * world * sizeof(KcMail) : 0;
}

extern "C" int b200al_kcenter_run_sharded(b200al_ctx* ctx, const float* d_emb_local, int n_local, int64_t row_base,
                                          int D, int metric, int criteria, int n_centres, void* d_state,
                                          uint8_t* d_taken, int budget, int64_t* d_picks, float* d_pick_vals,
                                          int world, int rank, int epoch, void* d_mail_local,
                                          void* const* d_mail_peers, const float* const* d_emb_peers,
                                          const int64_t* row_bases, int timeout_ms, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (budget == 0) return B200AL_OK;
  if (!d_state || !d_picks || n_local < 0 || budget < 0 || (n_local > 0 && (!d_emb_local || !d_taken)))
    return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (world < 1 || world > kKcMaxWorld || rank < 0 || rank >= world)
    return set_err(ctx, B200AL_ERR_ARG, "world %d / rank %d outside [1, %d]", world, rank, kKcMaxWorld);
  if (world > 1 && (!d_mail_local || !d_mail_peers || !d_emb_peers || !row_bases))
    return set_err(ctx, B200AL_ERR_ARG, "peer pointers missing");
  if (epoch < 1 || epoch > 30000) return set_err(ctx, B200AL_ERR_ARG, "epoch must be in [1, 30000]");
  if (budget > 65000) return set_err(ctx, B200AL_ERR_ARG, "budget above 65000 steps per call");
  if (metric != B200AL_L2 && metric != B200AL_COSINE) return set_err(ctx, B200AL_ERR_ARG, "bad metric");
  if (criteria != B200AL_MIN && criteria != B200AL_MEAN) return set_err(ctx, B200AL_ERR_ARG, "bad criteria");
  if (criteria == B200AL_MEAN && n_centres < 1) return set_err(ctx, B200AL_ERR_ARG, "mean criteria needs n_centres >= 1");
  if (int rc = check_dim(ctx, D)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  KcPersistParams p;
  memset(&p, 0, sizeof(p));
  p.emb = d_emb_local; p.n_local = n_local; p.row_base = row_base;
  p.metric = metric; p.criteria = criteria; p.n_centres = n_centres; p.budget = budget;
  p.state = d_state; p.taken = d_taken; p.picks = (long long*)d_picks; p.pick_vals = d_pick_vals;
  p.world = world; p.rank = rank; p.epoch = epoch;
  p.mail_local = (KcMail*)d_mail_local;
  if (world > 1) {
    for (int r = 0; r < world; ++r) {
      if (!d_mail_peers[r] || (!d_emb_peers[r] && row_bases[r + 1] > row_bases[r]))
        return set_err(ctx, B200AL_ERR_ARG, "peer pointer of rank %d is NULL", r);
      p.mail_peer[r] = (KcMail*)d_mail_peers[r];
      p.emb_peer[r] = d_emb_peers[r];
      p.base_peer[r] = row_bases[r];
    }
    p.base_peer[world] = row_bases[world];
    if (row_bases[rank] != row_base || row_bases[rank + 1] - row_bases[rank] != n_local)
      return set_err(ctx, B200AL_ERR_ARG, "row_bases do not match this rank's shard");
  }
  p.timeout_cycles = timeout_ms > 0 ? (long long)timeout_ms * 2000000LL : 0;
  if (int rc = run_kc_persist(ctx, D, p, st)) return rc;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

extern "C" int b200al_kcenter_run(b200al_ctx* ctx, const float* d_emb, int n, int D, int metric, int criteria,
                                  int n_centres, void* d_state, uint8_t* d_taken, int budget, int64_t* d_picks,
                                  float* d_pick_vals, void* stream) {
  if (!ctx) return set_err(nullptr, B200AL_ERR_ARG, "ctx is NULL");
  if (budget == 0) return B200AL_OK;
  if (!d_emb || !d_state || !d_taken || !d_picks || n < 1 || budget < 0) return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (metric != B200AL_L2 && metric != B200AL_COSINE) return set_err(ctx, B200AL_ERR_ARG, "bad metric");
  if (criteria != B200AL_MIN && criteria != B200AL_MEAN) return set_err(ctx, B200AL_ERR_ARG, "bad criteria");
  if (criteria == B200AL_MEAN && n_centres < 1) return set_err(ctx, B200AL_ERR_ARG, "mean criteria needs n_centres >= 1");
  if (int rc = check_dim(ctx, D)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (env_int("B200AL_KC_PERSIST", 1)) {
    // all `budget` steps in one cooperative launch (kcenter.cuh: kcenter_persist_kernel); same arithmetic, same picks
    KcPersistParams p;
    memset(&p, 0, sizeof(p));
    p.emb = d_emb; p.n_local = n; p.row_base = 0;
    p.metric = metric; p.criteria = criteria; p.n_centres = n_centres; p.budget = budget;
    p.state = d_state; p.taken = d_taken; p.picks = (long long*)d_picks; p.pick_vals = d_pick_vals;
    p.world = 1; p.rank = 0; p.epoch = 1;
    if (int rc = run_kc_persist(ctx, D, p, st)) return rc;
    CUDA_TRY(ctx, cudaGetLastError());
    return B200AL_OK;
  }
  int blocks = (n + 63) / 64;  // >= 8 rows per warp
  if (blocks > ctx->kc_blocks) blocks = ctx->kc_blocks;
  if (blocks < 1) blocks = 1;
  for (int step = 0; step < budget; ++step) {
    const int upd = step > 0;
    const double inv = 1.0 / (double)(n_centres + step);  // centres folded in once this step's update is done
    switch (D / 128) {
      case 1: launch_kc_step<1>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
      case 2: launch_kc_step<2>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
      case 4: launch_kc_step<4>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
      case 8: launch_kc_step<8>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
      case 12: launch_kc_step<12>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
      case 16: launch_kc_step<16>(d_emb, n, ctx->kc_centre, upd, metric, criteria, d_state, inv, d_taken, ctx->kc_part_val, ctx->kc_part_idx, blocks, st); break;
    }
    kcenter_pick_kernel<<<1, 256, 0, st>>>(ctx->kc_part_val, ctx->kc_part_idx, blocks, d_emb, D, d_taken,
                                          ctx->kc_centre, (long long*)d_picks, d_pick_vals, step);
    ctx->launches += 2;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// introspection / timing
// ---------------------------------------------------------------------------
extern "C" int b200al_kernel_launches(const b200al_ctx* ctx, int64_t* out_count) {
  if (!ctx || !out_count) return B200AL_ERR_ARG;
  *out_count = ctx->launches;
  return B200AL_OK;
}

extern "C" int b200al_set_timing(b200al_ctx* ctx, int enable) {
  if (!ctx) return B200AL_ERR_ARG;
  ctx->timing = enable != 0;
  ctx->ev_used = 0;
  ctx->conv_ms = 0;
  ctx->conv_flops = 0;
  ctx->conv_launches = 0;
  ctx->conv_timed = 0;
  memset(ctx->layer_ms, 0, sizeof(ctx->layer_ms));
  return B200AL_OK;
}

// Synchronises the device, folds recorded event pairs into the running total.
extern "C" int b200al_conv_time_ms(b200al_ctx* ctx, float* out_ms, int64_t* out_launches, double* out_flops) {
  if (!ctx || !out_ms) return B200AL_ERR_ARG;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaDeviceSynchronize());
  for (size_t i = 0; i + 1 < ctx->ev_used; i += 2) {
    float ms = 0.f;
    CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev[i], ctx->ev[i + 1]));
    ctx->conv_ms += ms;
    ctx->layer_ms[ctx->ev_layer[i / 2] & 31] += ms;
  }
  ctx->ev_used = 0;
  *out_ms = (float)ctx->conv_ms;
  if (out_launches) *out_launches = ctx->conv_timed;
  if (out_flops) *out_flops = ctx->conv_flops;
  return B200AL_OK;
}

// ---------------------------------------------------------------------------
// test hooks
// ---------------------------------------------------------------------------
extern "C" int b200al_debug_layer_times(b200al_ctx* ctx, float* out_ms, int n) {
  if (!ctx || !out_ms) return B200AL_ERR_ARG;
  for (int i = 0; i < n && i < 32; ++i) out_ms[i] = (float)ctx->layer_ms[i];
  return B200AL_OK;
}

// info: kind, cin, cout, stride, inH, inW, in_pitch, in_c_off, outH, outW, out_pitch, out_c_off
extern "C" int b200al_debug_layer_info(b200al_ctx* ctx, int layer_index, int* info) {
  if (!ctx || !info || layer_index < 0 || layer_index >= (int)ctx->layers.size()) return B200AL_ERR_ARG;
  const Layer& L = ctx->layers[layer_index];
  const ActBuf& ib = ctx->bufs[L.in.buf];
  const ActBuf& ob = ctx->bufs[L.out.buf];
  info[0] = L.kind; info[1] = L.cin; info[2] = L.cout; info[3] = L.stride;
  info[4] = ctx->H >> ib.level; info[5] = ctx->W >> ib.level; info[6] = 2 * ib.c_total; info[7] = L.in.c_off;
  info[8] = ctx->H >> ob.level; info[9] = ctx->W >> ob.level; info[10] = 2 * ob.c_total; info[11] = L.out.c_off;
  return B200AL_OK;
}

extern "C" int b200al_debug_layer(b200al_ctx* ctx, int layer_index, int impl, const void* d_in, void* d_out,
                                  int n_samples, const float* d_mask, void* stream) {
  if (!ctx || !d_in || !d_out || layer_index < 0 || layer_index >= (int)ctx->layers.size())
    return set_err(ctx, B200AL_ERR_ARG, "bad argument");
  if (!ctx->weights_loaded) return set_err(ctx, B200AL_ERR_STATE, "weights not loaded");
  if (n_samples < 1 || n_samples > ctx->max_samples) return set_err(ctx, B200AL_ERR_ARG, "bad n_samples");
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Layer& L = ctx->layers[layer_index];
  const ActBuf& ib = ctx->bufs[L.in.buf];
  const ActBuf& ob = ctx->bufs[L.out.buf];
  const int inH = ctx->H >> ib.level, inW = ctx->W >> ib.level;
  const float* mask = (d_mask && L.kind == KIND_CONV3) ? d_mask + L.mask_off : nullptr;
  if (impl == 0) {
    CUtensorMap tmA, tmO;
    int rc = make_layer_maps(ctx, L, (const __nv_bfloat16*)d_in, (const __nv_bfloat16*)d_out, n_samples, &tmA, &tmO);
    if (rc) return rc;
    ConvTcParams p = L.p;
    p.out = (__nv_bfloat16*)d_out;
    p.mask = mask;
    return launch_conv(ctx, L, tmA, tmO, p, n_samples, st);
  }
  ConvSimtParams q;
  q.in = (const __nv_bfloat16*)d_in + L.in.c_off;
  q.in_pitch = 2 * ib.c_total; q.in_lo_off = ib.c_total; q.in_H = inH; q.in_W = inW; q.cin = L.cin;
  q.wp = L.wp; q.cout = L.cout; q.stride = L.stride; q.mode = L.kind == KIND_CONV3 ? 0 : 1;
  q.use_lo = ctx->nsplit == 3;
  q.scale = L.scale; q.bias = L.bias; q.shift = L.shift; q.mask = mask; q.mask_stride = B200AL_MASK_CHANNELS;
  q.slope = L.p.slope;
  q.out = (__nv_bfloat16*)d_out;
  q.out_pitch = 2 * ob.c_total; q.out_hi_off = L.out.c_off; q.out_lo_off = ob.c_total + L.out.c_off;
  q.out_H = ctx->H >> ob.level; q.out_W = ctx->W >> ob.level; q.write_lo = ctx->nsplit == 3;
  conv_simt_kernel<<<ctx->sm_count * 16, 256, 0, st>>>(q, n_samples);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return B200AL_OK;
}
