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

#include <cooperative_groups.h>
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

// ---------------------------------------------------------------------------------------------------------------
// Persistent greedy loop: ALL `budget` steps of kcenter_greedy (coreset_selector.py:29-45) in ONE cooperative launch
// instead of two launches per step from the host.  Per step every block folds the newest centre into the running
// state of its own rows and publishes a (value, row) partial; one grid-wide barrier later every block reduces the
// partials redundantly (no second barrier), the owner of the winning row marks it taken, block 0 records the pick.
//
// Sharded across GPUs (one process per GPU, the rows cut into contiguous ranges): the same kernel runs on every rank
// over its own rows.  After the grid-wide reduction block 0 writes the rank's candidate (value, global row, tag) into
// a mailbox slot in EVERY peer's memory (NVLink peer stores, st.release.sys); thread 0 of every block polls the LOCAL
// mailboxes (ld.acquire.sys) until the step's W candidates have landed, takes the largest value / lowest global row,
// and the block reads the winner's embedding row straight from its owner's shard (a 2 KB peer load).  Nothing else
// crosses the links: 16 bytes per rank and step plus the chosen centre - the exchange of the north star's acquisition
// pass - and compute, exchange and the next step's update are one kernel.
// Mailboxes are double-buffered by step parity (a rank can only publish step s+2 after every peer has published step
// s+1, which they do after having consumed step s); consecutive calls alternate between two banks of slots and tag
// every entry with (call epoch, step), so a fast rank entering the next call cannot disturb a slow one.
// Arithmetic is exactly that of kcenter_step_kernel / kcenter_step_mean_kernel: identical picks.
struct __align__(16) KcMail {
  float val;
  int tag;        // epoch * 65536 + step + 1 (0 = never written)
  long long pos;  // global row, or -1: nothing left on that rank
};
constexpr int kKcMaxWorld = 8;

struct KcPersistParams {
  const float* emb;      // local rows [n_local][D]
  int n_local;
  long long row_base;    // global index of local row 0
  int metric, criteria, n_centres, budget;
  void* state;           // float [n_local] running min, or double [n_local] running sum
  uint8_t* taken;        // [n_local]
  float* part_val;       // [2][gridDim.x]
  long long* part_idx;   // [2][gridDim.x]
  long long* picks;      // [budget] global rows
  float* pick_vals;      // nullable
  int world, rank, epoch;
  KcMail* mail_local;                    // [4][world] (bank = epoch parity, slot = step parity), written by the peers
  KcMail* mail_peer[kKcMaxWorld];        // rank r's mail_local as mapped here (mail_peer[rank] == mail_local)
  const float* emb_peer[kKcMaxWorld];    // rank r's shard rows as mapped here (emb_peer[rank] == emb)
  long long base_peer[kKcMaxWorld + 1];  // global row range of rank r: [base_peer[r], base_peer[r+1])
  long long timeout_cycles;              // mailbox poll watchdog (0 = none)
};

__device__ __forceinline__ void st_release_sys_s32(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int NV>
__global__ void __launch_bounds__(256) kcenter_persist_kernel(const KcPersistParams p) {
  constexpr int D = NV * 128;
  constexpr long long kNone = 0x7fffffffffffffffLL;
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int rows_per_block = (p.n_local + gridDim.x - 1) / gridDim.x;
  const int r_begin = min(p.n_local, (int)blockIdx.x * rows_per_block);
  const int r_end = min(p.n_local, r_begin + rows_per_block);
  __shared__ float sv[32];
  __shared__ long long si[32];
  __shared__ float best_v;
  __shared__ long long best_i;
  float* mind = reinterpret_cast<float*>(p.state);
  double* msum = reinterpret_cast<double*>(p.state);
  const float* centre = nullptr;
  for (int step = 0; step < p.budget; ++step) {
    float4 c[NV];
    if (centre) load_row<NV>(centre, lane, c);
    const double inv = 1.0 / (double)(p.n_centres + step);
    float bv = -CUDART_INF_F;
    long long bi = kNone;
    // Two rows per iteration: both rows' loads are in flight before either reduction starts (a single row per warp
    // leaves the step latency-bound at ~half of the HBM rate: load -> fp64 FMA chain -> 5-stage fp64 shuffle tree).
    // Narrow rows only: 2 x NV float4 of x plus the centre must stay in registers.
    constexpr int RPI = NV <= 4 ? 2 : 1;
    for (int row0 = r_begin + warp; row0 < r_end; row0 += RPI * nw) {
      float4 x[RPI][NV];
      float d[RPI];
      if (centre) {
#pragma unroll
        for (int u = 0; u < RPI; ++u) {
          const int row = min(row0 + u * nw, r_end - 1);
          load_row<NV>(p.emb + (size_t)row * D, lane, x[u]);
        }
#pragma unroll
        for (int u = 0; u < RPI; ++u) d[u] = row_distance<NV>(x[u], c, p.metric);
      }
#pragma unroll
      for (int u = 0; u < RPI; ++u) {  // ascending rows: row0, row0 + nw
        const int row = row0 + u * nw;
        if (row >= r_end) break;
        float v;
        if (p.criteria == CRITERIA_MEAN) {
          double m = msum[row];
          if (centre) {
            m += (double)d[u];
            if (lane == 0) msum[row] = m;
          }
          v = (float)(m * inv);
        } else {
          v = mind[row];
          if (centre) {
            v = fminf(v, d[u]);
            if (lane == 0) mind[row] = v;
          }
        }
        if (!p.taken[row] && v > bv) { bv = v; bi = p.row_base + row; }  // rows ascend: strict > keeps the first max
      }
    }
    if (lane == 0) { sv[warp] = bv; si[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < nw; ++w)
        if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
      p.part_val[(step & 1) * gridDim.x + blockIdx.x] = bv;
      p.part_idx[(step & 1) * gridDim.x + blockIdx.x] = bi;
    }
    grid.sync();
    // every block: the same reduction over all partials (written by other SMs: read at L2)
    bv = -CUDART_INF_F;
    bi = kNone;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
      const float v = __ldcg(p.part_val + (step & 1) * gridDim.x + i);
      const long long ix = __ldcg(p.part_idx + (step & 1) * gridDim.x + i);
      if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float v = __shfl_xor_sync(0xffffffffu, bv, o);
      const long long ix = __shfl_xor_sync(0xffffffffu, bi, o);
      if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
    }
    if (lane == 0) { sv[warp] = bv; si[warp] = bi; }  // (the block reduce above is done with sv / si: grid.sync)
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < nw; ++w)
        if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
      if (bi == kNone) bi = -1;
      if (p.world > 1) {
        const int tag = p.epoch * 65536 + step + 1;
        const int slot = (((p.epoch & 1) << 1) | (step & 1)) * p.world;
        if (blockIdx.x == 0) {  // publish this rank's candidate into every rank's mailbox (its own included)
          for (int r = 0; r < p.world; ++r) {
            KcMail* m = p.mail_peer[r] + slot + p.rank;
            m->val = bv;
            m->pos = bi;
            st_release_sys_s32(&m->tag, tag);
          }
        }
        const long long t0 = clock64();
        float gv = -CUDART_INF_F;
        long long gi = -1;
        for (int r = 0; r < p.world; ++r) {
          const KcMail* m = p.mail_local + slot + r;
          while (ld_acquire_sys_s32(&m->tag) != tag) {
            if (p.timeout_cycles && clock64() - t0 > p.timeout_cycles) {
              printf("b200al: k-center rank %d timed out waiting for rank %d at step %d\n", p.rank, r, step);
              __trap();
            }
          }
          const float v = *reinterpret_cast<const volatile float*>(&m->val);
          const long long ix = *reinterpret_cast<const volatile long long*>(&m->pos);
          if (ix >= 0 && (gi < 0 || v > gv || (v == gv && ix < gi))) { gv = v; gi = ix; }
        }
        bv = gv;
        bi = gi;
      }
      best_v = bv;
      best_i = bi;
    }
    __syncthreads();
    const long long best = best_i;
    if (best < 0) {  // nothing left anywhere: the same decision in every block of every rank
      if (blockIdx.x == 0 && threadIdx.x == 0)
        for (int s2 = step; s2 < p.budget; ++s2) { p.picks[s2] = -1; if (p.pick_vals) p.pick_vals[s2] = -CUDART_INF_F; }
      return;
    }
    const long long local = best - p.row_base;
    if (threadIdx.x == 0) {
      if (local >= r_begin && local < r_end) p.taken[local] = 1;
      if (blockIdx.x == 0) { p.picks[step] = best; if (p.pick_vals) p.pick_vals[step] = best_v; }
    }
    // next centre: the winner's row, from this rank's rows or straight from its owner's shard over NVLink
    if (p.world > 1) {
      int owner = 0;
      while (owner + 1 < p.world && best >= p.base_peer[owner + 1]) ++owner;
      centre = p.emb_peer[owner] + (size_t)(best - p.base_peer[owner]) * D;
    } else {
      centre = p.emb + (size_t)local * D;
    }
    __syncthreads();  // taken[] / best_i are settled before the next step's row loop and reduction
  }
}

}  // namespace b200al
