/* libb200al - B200-native active-learning acquisition pass (C ABI).
 *
 * Drop-in boundary for the hot path of trnKhanh/medical-image-analysis:
 * `ActiveSelector.select_next_batch` (src/activelearning/active_selector.py:10-19)
 * and what it runs on a pool: the UNet forward, the per-image uncertainty
 * reduction, rank & cut, and k-center-greedy on bottleneck embeddings.  The
 * reference is pure Python with no FFI, so these entry points
 * are what a ctypes binding on the reference side calls (INTEGRATION.md shows
 * the stub).  Each one cites the reference code it replaces.
 *
 * Conventions: every pointer named d_* is a DEVICE pointer owned by the caller
 * (e.g. torch-allocated); the library owns only what hangs off the opaque
 * context.  Calls are asynchronous on `stream` (a cudaStream_t passed as
 * void*; NULL = legacy default stream) unless stated.  Return value 0 on
 * success, a negative b200al_status otherwise; b200al_last_error() gives the
 * message.  There is no CPU fallback anywhere: without a sm_100 GPU the
 * compute calls fail with B200AL_ERR_CUDA.
 */
#ifndef B200AL_H_
#define B200AL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200al_ctx b200al_ctx;

enum b200al_status {
  B200AL_OK = 0,
  B200AL_ERR_ARG = -1,         /* bad argument / unsupported shape */
  B200AL_ERR_CUDA = -2,        /* CUDA runtime / driver error */
  B200AL_ERR_STATE = -3,       /* call order (e.g. weights not loaded) */
  B200AL_ERR_UNSUPPORTED = -4  /* configuration outside the al_train UNet */
};

enum b200al_score_mode {    /* which selector's score */
  B200AL_ENTROPY = 0,       /* EntropySelector.cal_scores     entropy_selector.py:45-49 */
  B200AL_MARGIN = 1,        /* MarginSelector.cal_scores      margin_selector.py:45-48 */
  B200AL_CONFIDENCE = 2,    /* ConfidenceSelector.cal_scores  confidence_selector.py:45-47 */
  B200AL_DICE = 3           /* dice-uncertainty (not a reference selector; BASELINE north star):
                               DiceLoss(pbar, one_hot(argmax pbar)) with the semantics of losses/dice_loss.py:32-72
                               (do_bg = True, batch = False, squared = False); the `smooth` argument of the scoring
                               calls is DiceLoss.smooth (1e-5) in this mode */
};
enum b200al_metric { B200AL_L2 = 0, B200AL_COSINE = 1 }; /* CoresetSelector(metric=...) */
enum b200al_criteria { B200AL_MIN = 0, B200AL_MEAN = 1 }; /* CoresetSelector(coreset_criteria=...) */
enum b200al_precision {
  B200AL_BF16X3 = 0, /* bf16 hi/lo split operands, 3 MMAs per k-step, fp32 accumulate (default; parity grade) */
  B200AL_BF16 = 1,   /* single bf16 operands (fast; does not meet the 2e-2 logit tolerance) */
  /* Per-layer mix for the SCORING path (b200al_unet_scores): the deep layers - encoder.levels.2.1 ... decoder.upsamples.2,
   * 64x64 resolution and below - contract single fp16 operands (one MMA per product, one activation plane), everything
   * else stays bf16x3.  Emulated on the golden probes (tools/precision_table.py, profiles/r02_precision.json): 7e-3 max
   * logit error, 8e-6 relative score error against the fp32 reference (tolerances 2e-2 / 1e-3).  Embeddings
   * (b200al_unet_embed), BADGE and the test hooks run bf16x3 throughout in this mode. */
  B200AL_MIXED = 2,
  B200AL_MIXED_DEEP = 3 /* the same with the smaller group encoder.levels.3.1 ... decoder.upsamples.1 (2.5e-3 / 3e-6) */
};
enum b200al_image_dtype { B200AL_U8 = 0 /* x/255, torchvision to_tensor */, B200AL_F32 = 1 };

#define B200AL_NUM_BLOCKS 18       /* PlainBlocks: 10 encoder + 8 decoder (unet.py:53-78,131-176) */
#define B200AL_NUM_UPSAMPLES 4     /* ConvTranspose2d k2 s2 (unet.py:143-144) */
#define B200AL_MASK_CHANNELS 2944  /* sum of Dropout2d channel counts over the 18 blocks */
#define B200AL_EMBED_DIM 512

/* One PlainBlock (blocks.py:83-102): Conv2d 3x3 (weight [Cout,Cin,3,3], bias) +
 * BatchNorm2d running statistics and affine.  fp32 device pointers in the
 * reference's state_dict layout (`...all.0.weight`, `...all.2.running_var`, ...). */
typedef struct {
  const float* d_conv_weight;
  const float* d_conv_bias;
  const float* d_bn_gamma;
  const float* d_bn_beta;
  const float* d_bn_mean;
  const float* d_bn_var;
} b200al_block_weights;

typedef struct {
  b200al_block_weights blocks[B200AL_NUM_BLOCKS];   /* forward order: encoder.levels.l.{0,1}, decoder.levels.l.{0,1} */
  const float* d_up_weight[B200AL_NUM_UPSAMPLES];   /* decoder.upsamples.l.weight [Cin,Cout,2,2] */
  const float* d_up_bias[B200AL_NUM_UPSAMPLES];
  const float* d_head_weight;                       /* decoder.seg_output.weight [K,32,1,1] */
  const float* d_head_bias;                         /* [K] */
  float bn_eps;                                     /* 1e-5  (blocks.py:98) */
  float leaky_slope;                                /* 0.01  (blocks.py:100) */
} b200al_weights;

/* Context for one device.  num_out = output classes K (num_classes+1,
 * al_trainer.py:508), in_channels must be 1, H = W = multiple of 128
 * (256 after JointResize, entry/activelearning/train.py:30), max_samples =
 * largest number of (image x MC repeat) samples forwarded per call chunk. */
int b200al_ctx_create(b200al_ctx** out, int device, int num_out, int in_channels, int H, int W, int max_samples,
                      int precision);
void b200al_ctx_destroy(b200al_ctx* ctx);
const char* b200al_last_error(const b200al_ctx* ctx); /* ctx may be NULL: last global error */

/* Fold BatchNorm and pack weights for the tensor-core kernels, on device.
 * Replaces reading `model.state_dict()` each round (al_trainer.py:1048-1083). */
int b200al_load_weights(b200al_ctx* ctx, const b200al_weights* w, void* stream);

/* Pool-image input pipeline (the dataset side of the reference: FUGCDataset.get_sample fugc_dataset.py:140-160 /
 * BUSIDataset.get_sample busi_dataset.py:126-143 -> F.to_tensor -> JointResize joint_transform.py:20-25 ->
 * ZScoreNormalize normalization.py:13-21) for n grayscale uint8 images of ragged sizes held in one device buffer:
 * image i = d_raw + d_offsets[i], d_hw[2i] rows x d_hw[2i+1] columns, tightly packed.  max_h / max_w bound the
 * source sizes (they size the antialias filter tables; down-scale factors up to 17 are supported).
 * d_images_out [n,H,W] fp32 = u8/255 resized with torch's antialiased bilinear filter.
 * d_norm_out nullable [n][2] = per-image (mean, unbiased std clipped at 1e-8) of the resized image; hand it to
 * b200al_unet_scores / b200al_unet_embed as d_norm to apply ZScoreNormalize inside the first conv's loader. */
int b200al_prep_images(b200al_ctx* ctx, const uint8_t* d_raw, const int64_t* d_offsets, const int32_t* d_hw, int n,
                       int max_h, int max_w, float* d_images_out, float* d_norm_out, void* stream);

/* Score n pool images: UNet.forward (unet.py:293) + the selector's per-image
 * score, T >= 1 MC-dropout repeats (T == 1 and drop_p == 0: plain eval mode,
 * exactly the reference selectors).  d_images: [n,H,W] u8 or f32.
 * d_norm nullable [n][2]: per-image (mean, std); pixels become (x - mean) / std on load (ZScoreNormalize,
 * normalization.py:21; --do-normalize), zero padding stays zero.
 * MC masks: d_drop_masks ([n,T,2944] multipliers, 0 or 1/(1-p)) if non-NULL,
 * else generated from (mc_seed, base_pos + i, t, channel) - see oracle/synth.py.
 * d_scores_out [n] fp32.  d_logits_out nullable: [n*T, K, H, W] fp32.
 * n*T must be <= max_samples. */
int b200al_unet_scores(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n, int mode,
                       int T, float drop_p, uint64_t mc_seed, int64_t base_pos, const float* d_drop_masks,
                       float smooth, float* d_scores_out, float* d_logits_out, void* stream);

/* UNet.get_enc_feature (unet.py:296-297, 89-93): encoder + global average pool.
 * d_emb_out [n,512] fp32.  n <= max_samples. */
int b200al_unet_embed(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n,
                      float* d_emb_out, void* stream);

/* BADGESelector.cal_scores (badge_selector.py:56-104): per image the gradient of
 * CrossEntropyLoss(logits, preds) + DiceLoss(logits, preds) w.r.t. decoder.seg_output.weight, preds = argmax softmax
 * (:79-80, 91-95), in closed form over the last feature map instead of an autograd backward (csrc/badge.cuh).
 * dice_smooth / dice_do_bg: DiceLoss(smooth, do_bg) as the trainer configures it (al_trainer.py:785-791: 1e-5, True;
 * softmax=True, batch=False, squared=False are assumed).  d_emb_out [n][out_stride] fp32: K*32 entries in
 * `weight.grad.flatten()` order, zero-padded to out_stride (use 128 to feed b200al_kmeans_pp).  K <= 4. */
int b200al_badge_embed(b200al_ctx* ctx, const void* d_images, int image_dtype, const float* d_norm, int n,
                       float dice_smooth, int dice_do_bg, float* d_emb_out, int out_stride, void* stream);

/* Kernel 3 alone: scores from fp32 logits [n,T,K,HW] already in HBM
 * (softmax -> mean over T -> entropy / margin / confidence -> mean over pixels). */
int b200al_uncertainty_from_logits(b200al_ctx* ctx, const float* d_logits, int n, int T, int K, int HW, int mode,
                                   float smooth, float* d_scores_out, void* stream);

/* Scratch of b200al_topk and b200al_kcenter_fused lives in a workspace owned by the context.  It is allocated ONLY by
 * b200al_workspace_reserve (grows to at least `bytes`; synchronises the device when it reallocates) - the hot calls
 * themselves never allocate and return B200AL_ERR_STATE when the workspace is too small.  The *_bytes functions give
 * the size a call needs (no reference counterpart: the reference lets torch / numpy allocate as they go). */
size_t b200al_workspace_bytes_topk(int n);
size_t b200al_workspace_bytes_kcenter_fused(int n, int D1, int D2, int metric);
int b200al_workspace_reserve(b200al_ctx* ctx, size_t bytes);

/* Rank & cut (entropy_selector.py:76-82): k best of n scores, descending,
 * ties -> lower position first.  Writes k' = min(k,n) entries; positions are
 * base_pos + index.  d_out_scores [k'], d_out_pos [k'] int64.
 * Needs b200al_workspace_bytes_topk(n) bytes of reserved workspace. */
int b200al_topk(b200al_ctx* ctx, const float* d_scores, int n, int k, int64_t base_pos, float* d_out_scores,
                int64_t* d_out_pos, void* stream);

/* k-center greedy (coreset_selector.py:19-51).  criteria B200AL_MIN (":34", d_state = float [n] running minimum
 * distance to the centres) or B200AL_MEAN (":36", d_state = double [n] running SUM of distances; the mean's
 * divisor is the same for every row).
 * init: folds the L initial centres (the labeled rows, coreset_selector.py:85,226) into d_state;
 * accumulate != 0 continues from the existing state.
 * For COSINE both matrices must hold rows normalised by b200al_normalize_rows. */
int b200al_kcenter_init(b200al_ctx* ctx, const float* d_emb, int n, int D, const float* d_centres, int L, int metric,
                        int criteria, void* d_state, int accumulate, void* stream);
/* run: `budget` greedy steps over rows not marked in d_taken (uint8 [n]); n_centres = number of centres already
 * folded into d_state (the divisor of the mean).
 * d_picks [budget] int64 row indices in selection order (-1 when exhausted),
 * d_pick_vals nullable [budget] = the criterion value of each pick (max-min distance / max mean distance).
 * d_state / d_taken are updated in place. */
int b200al_kcenter_run(b200al_ctx* ctx, const float* d_emb, int n, int D, int metric, int criteria, int n_centres,
                       void* d_state, uint8_t* d_taken, int budget, int64_t* d_picks, float* d_pick_vals, void* stream);
/* The same greedy loop with the rows SHARDED over `world` GPUs (one process per GPU; no reference counterpart - the
 * reference is single-process): this rank owns rows [row_bases[rank], row_bases[rank+1]) = d_emb_local [n_local, D] with
 * its own d_state / d_taken, and ONE kernel per rank runs all `budget` steps: per step it updates its rows, publishes
 * its best (value, global row) as 16 bytes into every peer's mailbox over NVLink, waits for the peers' entries, and
 * reads the winning row straight from its owner's shard.  No other data crosses the links, no collective is called.
 * d_mail_local: this rank's mailbox (b200al_kcenter_mailbox_bytes(world) bytes, zeroed once when it is allocated);
 * d_mail_peers[r] / d_emb_peers[r]: rank r's mailbox / shard rows as mapped into THIS process (peer-accessible memory,
 * e.g. torch.distributed._symmetric_memory buffer_ptrs or cudaIpcOpenMemHandle; entry `rank` = the local pointers).
 * epoch: call counter starting at 1, the same on every rank, incremented per call on the same mailboxes.
 * timeout_ms > 0: a rank that waits longer for a peer traps instead of hanging the GPU.  Every rank gets the same
 * d_picks (GLOBAL row indices, selection order).  world == 1 needs no peer pointers (what b200al_kcenter_run calls). */
size_t b200al_kcenter_mailbox_bytes(int world);
int b200al_kcenter_run_sharded(b200al_ctx* ctx, const float* d_emb_local, int n_local, int64_t row_base, int D,
                               int metric, int criteria, int n_centres, void* d_state, uint8_t* d_taken, int budget,
                               int64_t* d_picks, float* d_pick_vals, int world, int rank, int epoch,
                               void* d_mail_local, void* const* d_mail_peers, const float* const* d_emb_peers,
                               const int64_t* row_bases, int timeout_ms, void* stream);
/* CoresetSelector with externally loaded features, coreset_fusion="add" (coreset_selector.py:110-157, 221-229):
 * k-center greedy on  loaded_weight * D_loaded / sum(D_loaded) + emb_weight * D_emb / sum(D_emb)  (emb_weight =
 * fl32(1 - w) computed by the caller in double like the reference's python expression) over rows 0..n-1 of the two
 * feature matrices d_emb [n,D1] (D1 <= 1024) and d_loaded [n,D2] (any width); rows 0..n_init-1 are the labeled set.
 * The two global pair-distance sums are computed on the device (one O(n^2) pass each, no n x n matrix).
 * d_picks [budget] int64 in selection order (-1 when exhausted), d_pick_vals nullable.  Asynchronous on `stream`;
 * needs b200al_workspace_bytes_kcenter_fused(n, D1, D2, metric) bytes of reserved workspace. */
int b200al_kcenter_fused(b200al_ctx* ctx, const float* d_emb, int D1, const float* d_loaded, int D2, int n, int n_init,
                         int metric, int criteria, float loaded_weight, float emb_weight, int budget,
                         int64_t* d_picks, float* d_pick_vals, void* stream);

/* KMeanSelector (kmean_selector.py:19-196).
 * standardize_rows: (x - mean) / std per row, population std (:95-97).
 * kmeans_pp: scikit-learn's weighted k-means++ seeding (`kmeans_plusplus`, kmean_selector.py:190-192; sklearn 1.9.0
 * `_kmeans_plusplus`) on d_X [n,D] fp32 with normalised sample weights d_weight [n] fp32.  The caller draws the random
 * numbers from numpy's global RandomState in sklearn's order: first_centre = choice(n, p=w/w.sum()), then
 * d_uniform [(n_clusters-1) * n_local_trials] doubles = uniform(size=n_local_trials) per further centre
 * (n_local_trials = 2 + int(log(n_clusters)), at most 16).  d_scratch: b200al_kmeans_pp_scratch_bytes(n, trials)
 * bytes of device memory.  d_picks [n_clusters] int64 row indices in pick order. */
int b200al_standardize_rows(b200al_ctx* ctx, const float* d_in, float* d_out, int n, int D, void* stream);
size_t b200al_kmeans_pp_scratch_bytes(int n, int n_local_trials);
int b200al_kmeans_pp(b200al_ctx* ctx, const float* d_X, int n, int D, const float* d_weight, int first_centre,
                     const double* d_uniform, int n_clusters, int n_local_trials, void* d_scratch, int64_t* d_picks,
                     void* stream);
int b200al_normalize_rows(b200al_ctx* ctx, const float* d_in, float* d_out, int n, int D, void* stream);

/* Rank & cut WITHOUT any allocation - no counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same result as b200al_topk (k best of n scores, descending, ties -> lower position first), but it needs no
 * context and its scratch (64-bit keys, candidates, radix state) lives in a file-scope `static __device__` array of
 * the library, sized for n <= B200AL_TOPK_STATIC_MAX_N scores (4 MiB + 1 KiB, held by every GPU that loads the
 * library until the process exits).  n beyond that is rejected with B200AL_ERR_ARG.  `device` is made current; the
 * first call on a device resolves the array's device address (cudaGetSymbolAddress) - make that call outside a
 * graph capture.  No cudaMalloc / cudaFree / allocating library routine is called, ever.
 * ALL calls on one GPU share the one array: it keeps the previous call's contents (everything that must start at
 * zero is cleared by the call itself), and two calls that overlap on the GPU (different streams, or different host
 * threads) race - serialise them.  Asynchronous on `stream`. */
#define B200AL_TOPK_STATIC_MAX_N 262144
int b200al_topk_static(int device, const float* d_scores, int n, int k, int64_t base_pos, float* d_out_scores,
                       int64_t* d_out_pos, void* stream);

/* Introspection for benches / tests. */
int b200al_kernel_launches(const b200al_ctx* ctx, int64_t* out_count);      /* kernels launched so far */
/* CUDA-event timing around every tcgen05 conv launch (up to 32768 launches
 * between two b200al_conv_time_ms calls).  conv_time_ms synchronises the
 * device and returns the summed kernel time, the number of timed launches and
 * their algorithmic FLOPs (2*MAC, single-precision-equivalent). */
int b200al_set_timing(b200al_ctx* ctx, int enable);
int b200al_conv_time_ms(b200al_ctx* ctx, float* out_ms, int64_t* out_launches, double* out_flops);

/* TEST HOOKS (not part of the drop-in surface): run ONE 3x3 conv / transposed
 * conv layer of the loaded model on caller-provided NHWC hi/lo activations,
 * with the tcgen05 kernel (impl 0) or the SIMT cross-check kernel (impl 1). */
int b200al_debug_layer(b200al_ctx* ctx, int layer_index, int impl, const void* d_in, void* d_out, int n_samples,
                       const float* d_mask, void* stream);
int b200al_debug_layer_info(b200al_ctx* ctx, int layer_index, int* info /*[12]*/);
/* per-layer sums of the times folded by the last b200al_conv_time_ms call(s) since b200al_set_timing */
int b200al_debug_layer_times(b200al_ctx* ctx, float* out_ms, int n);

#ifdef __cplusplus
}
#endif
#endif /* B200AL_H_ */
