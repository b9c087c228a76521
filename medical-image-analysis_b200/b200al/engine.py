"""Device-side engine: owns a libb200al context, ingests UNet weights from a
live `nn.Module.state_dict()` (reference key layout) and runs the
scoring / embedding / rank / k-center calls.  torch is used for device memory
and streams only; all arithmetic happens in the CUDA library.
"""
from __future__ import annotations

import ctypes as C
from typing import Mapping

import torch

from . import _lib

CHANNELS = (32, 64, 128, 256, 512)


def block_prefixes():
    """state_dict prefixes of the 18 PlainBlocks in forward order
    (src/models/unet/unet.py:53-78,131-176)."""
    out = []
    for l in range(5):
        out += [f"encoder.levels.{l}.0.all", f"encoder.levels.{l}.1.all"]
    for l in range(4):
        out += [f"decoder.levels.{l}.0.all", f"decoder.levels.{l}.1.all"]
    return out


def block_channels():
    c = []
    for l in range(5):
        c += [CHANNELS[l], CHANNELS[l]]
    for l in range(4):
        c += [CHANNELS[3 - l], CHANNELS[3 - l]]
    return c


def validate_state_dict(sd: Mapping[str, torch.Tensor]) -> int:
    """Checks that `sd` is the al_train UNet (2D, plain blocks, BatchNorm,
    transposed-conv upsampling, no deep supervision, 1 input channel) and
    returns K.  Anything else raises NotImplementedError (no silent fallback)."""
    if any(".downsample_skip." in k for k in sd):
        raise NotImplementedError("residual blocks (block_type='res') are not supported")
    if any(k.startswith("decoder.ds.") for k in sd):
        raise NotImplementedError("deep-supervision heads are not supported")
    pre = block_prefixes()
    chans = block_channels()
    cin = [1] + [chans[i - 1] for i in range(1, 10)] + [0] * 8
    for l in range(4):
        cin[10 + 2 * l] = 2 * CHANNELS[3 - l]
        cin[11 + 2 * l] = CHANNELS[3 - l]
    for i, p in enumerate(pre):
        w = sd.get(f"{p}.0.weight")
        if w is None:
            raise NotImplementedError(f"missing {p}.0.weight: not the al_train UNet ([32,64,128,256,512], plain)")
        if w.dim() != 4:
            raise NotImplementedError("3D UNet is not supported")
        if f"{p}.2.running_mean" not in sd:
            raise NotImplementedError("normalization must be 'batch' (InstanceNorm has no running statistics to fold)")
        if tuple(w.shape) != (chans[i], cin[i], 3, 3):
            if i == 0 and w.shape[1] != 1:
                raise NotImplementedError(f"in_channels={w.shape[1]}: only grayscale (1) is supported")
            raise NotImplementedError(f"{p}.0.weight has shape {tuple(w.shape)}, expected {(chans[i], cin[i], 3, 3)}")
    for l in range(4):
        w = sd.get(f"decoder.upsamples.{l}.weight")
        if w is None or tuple(w.shape) != (CHANNELS[4 - l], CHANNELS[3 - l], 2, 2):
            raise NotImplementedError("upsampling must be ConvTranspose2d(k=2, s=2) (upconv=True)")
    hw = sd.get("decoder.seg_output.weight")
    if hw is None or hw.dim() != 4 or hw.shape[1] != 32 or hw.shape[2:] != (1, 1):
        raise NotImplementedError("decoder.seg_output must be a 1x1 conv on 32 channels")
    return int(hw.shape[0])


def module_hparams(model) -> tuple[float, float, float | None]:
    """(bn_eps, leaky_slope, dropout_p) read from the live module when it is
    one; the reference's defaults otherwise (blocks.py:98,100)."""
    eps, slope, p = 1e-5, 0.01, None
    mods = getattr(model, "modules", None)
    if callable(mods):
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                eps = float(m.eps)
            elif isinstance(m, torch.nn.LeakyReLU):
                slope = float(m.negative_slope)
            elif isinstance(m, torch.nn.Dropout2d) and p is None:
                p = float(m.p)
    return eps, slope, p


class Engine:
    """One context per (device, K, H, W, max_samples, precision)."""

    def __init__(self, device, num_out: int, H: int = 256, W: int = 256, max_samples: int = 64,
                 precision: str = "mixed"):
        self.lib = _lib.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("b200al runs on a CUDA (sm_100a) device only; there is no CPU path")
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.K, self.H, self.W, self.max_samples = num_out, H, W, max_samples
        self.precision = precision
        ctx = C.c_void_p()
        _lib.check(self.lib.b200al_ctx_create(C.byref(ctx), idx, num_out, 1, H, W, max_samples,
                                              _lib.PRECISIONS[precision]))
        self.ctx = ctx
        self._keep = []          # device tensors the loaded weights point at
        self.dropout_p = None

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.b200al_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ utils
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)

    def _check(self, rc):
        _lib.check(rc, self.ctx)

    # ---------------------------------------------------------------- weights
    def load_model(self, model, state_dict=None) -> None:
        """Fold + pack the weights of `model` (nn.Module with the reference
        UNet's state_dict layout, or a plain state_dict).  `state_dict` overrides
        the tensors (e.g. the copy broadcast from rank 0); hyper-parameters still
        come from `model`."""
        sd = state_dict if state_dict is not None else (model.state_dict() if hasattr(model, "state_dict") else model)
        K = validate_state_dict(sd)
        if K != self.K:
            raise ValueError(f"model has {K} output classes, engine was created for {self.K}")
        eps, slope, p = module_hparams(model)
        self.dropout_p = p

        def dev(name):
            t = sd[name].detach().to(device=self.device, dtype=torch.float32).contiguous()
            self._keep.append(t)
            return C.c_void_p(t.data_ptr())

        self._keep = []
        w = _lib.Weights()
        for i, pre in enumerate(block_prefixes()):
            b = w.blocks[i]
            b.d_conv_weight = dev(f"{pre}.0.weight")
            b.d_conv_bias = dev(f"{pre}.0.bias")
            b.d_bn_gamma = dev(f"{pre}.2.weight")
            b.d_bn_beta = dev(f"{pre}.2.bias")
            b.d_bn_mean = dev(f"{pre}.2.running_mean")
            b.d_bn_var = dev(f"{pre}.2.running_var")
        for l in range(4):
            w.d_up_weight[l] = dev(f"decoder.upsamples.{l}.weight")
            w.d_up_bias[l] = dev(f"decoder.upsamples.{l}.bias")
        w.d_head_weight = dev("decoder.seg_output.weight")
        w.d_head_bias = dev("decoder.seg_output.bias")
        w.bn_eps = eps
        w.leaky_slope = slope
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_load_weights(self.ctx, C.byref(w), self._stream()))

    # ---------------------------------------------------------------- scoring
    def _images(self, images: torch.Tensor):
        if images.device != self.device:
            raise ValueError(f"images must live on {self.device}")
        if images.dim() == 4:
            if images.shape[1] != 1:
                raise NotImplementedError("only 1-channel images are supported")
            images = images[:, 0]
        if images.shape[-2:] != (self.H, self.W):
            raise ValueError(f"images must be {self.H}x{self.W}, got {tuple(images.shape[-2:])}")
        if images.dtype == torch.uint8:
            return images.contiguous(), _lib.U8
        if images.dtype == torch.float32:
            return images.contiguous(), _lib.F32
        raise TypeError(f"image dtype {images.dtype} not supported (uint8 or float32)")

    def _norm(self, norm, n: int):
        if norm is None:
            return None
        norm = norm.to(device=self.device, dtype=torch.float32).contiguous()
        if tuple(norm.shape) != (n, 2):
            raise ValueError("norm must be [n,2] = per-image (mean, std)")
        return norm

    def prep(self, raw: torch.Tensor, offsets: torch.Tensor, hw: torch.Tensor, max_h: int, max_w: int,
             normalize: bool = False):
        """Dataset-side pipeline on the device (to_tensor -> JointResize -> ZScoreNormalize statistics) for n ragged
        grayscale uint8 images packed in `raw` (uint8 [bytes]); offsets int64 [n], hw int32 [n,2], all on device.
        Returns (images fp32 [n,H,W], norm [n,2] or None); pass `norm` on to scores / embed."""
        n = int(offsets.shape[0])
        for t, dt in ((raw, torch.uint8), (offsets, torch.int64), (hw, torch.int32)):
            if t.device != self.device or t.dtype != dt or not t.is_contiguous():
                raise ValueError("prep: raw uint8 / offsets int64 / hw int32 must be contiguous on the engine's device")
        if tuple(hw.shape) != (n, 2):
            raise ValueError("hw must be [n,2]")
        out = torch.empty((n, self.H, self.W), dtype=torch.float32, device=self.device)
        norm = torch.empty((n, 2), dtype=torch.float32, device=self.device) if normalize else None
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_prep_images(self.ctx, self._ptr(raw), self._ptr(offsets), self._ptr(hw), n,
                                                    int(max_h), int(max_w), self._ptr(out), self._ptr(norm),
                                                    self._stream()))
        return out, norm

    def scores(self, images: torch.Tensor, mode: str = "entropy", T: int = 1, drop_p: float = 0.0, seed: int = 0,
               base_pos: int = 0, masks: torch.Tensor | None = None, smooth: float = 1e-8,
               return_logits: bool = False, norm: torch.Tensor | None = None):
        """Per-image scores [n] (fp32, on device) for device-resident images
        [n,H,W] / [n,1,H,W]; chunks internally to max_samples // T images.
        norm: optional [n,2] per-image (mean, std) -> ZScoreNormalize applied while loading."""
        imgs, dt = self._images(images)
        n = imgs.shape[0]
        norm = self._norm(norm, n)
        out = torch.empty(n, dtype=torch.float32, device=self.device)
        logits = torch.empty((n * T, self.K, self.H, self.W), dtype=torch.float32, device=self.device) \
            if return_logits else None
        per = max(1, self.max_samples // T)
        if T > self.max_samples:
            raise ValueError(f"T={T} exceeds max_samples={self.max_samples}")
        if masks is not None:
            masks = masks.to(device=self.device, dtype=torch.float32).contiguous()
            if tuple(masks.shape) != (n, T, _lib.MASK_CHANNELS):
                raise ValueError(f"masks must be [n,T,{_lib.MASK_CHANNELS}]")
        st = self._stream()
        with torch.cuda.device(self.device):
            for i in range(0, n, per):
                m = min(per, n - i)
                self._check(self.lib.b200al_unet_scores(
                    self.ctx, self._ptr(imgs[i:i + m]), dt, self._ptr(norm[i:i + m] if norm is not None else None), m,
                    _lib.MODES[mode], T, float(drop_p), int(seed),
                    int(base_pos + i), self._ptr(masks[i:i + m] if masks is not None else None), float(smooth),
                    self._ptr(out[i:i + m]), self._ptr(logits[i * T:(i + m) * T] if logits is not None else None), st))
        return (out, logits) if return_logits else out

    def embed(self, images: torch.Tensor, norm: torch.Tensor | None = None) -> torch.Tensor:
        """UNet.get_enc_feature: [n,512] fp32 on device."""
        imgs, dt = self._images(images)
        n = imgs.shape[0]
        norm = self._norm(norm, n)
        out = torch.empty((n, _lib.EMBED_DIM), dtype=torch.float32, device=self.device)
        st = self._stream()
        with torch.cuda.device(self.device):
            for i in range(0, n, self.max_samples):
                m = min(self.max_samples, n - i)
                self._check(self.lib.b200al_unet_embed(self.ctx, self._ptr(imgs[i:i + m]), dt,
                                                       self._ptr(norm[i:i + m] if norm is not None else None), m,
                                                       self._ptr(out[i:i + m]), st))
        return out

    def badge_embed(self, images: torch.Tensor, norm: torch.Tensor | None = None, dice_smooth: float = 1e-5,
                    dice_do_bg: bool = True, pad_to: int = 128) -> torch.Tensor:
        """BADGE gradient embeddings (badge_selector.py:56-104): [n, pad_to] fp32 on device, the first K*32 columns
        are `decoder.seg_output.weight.grad.flatten()` of CE + Dice against the model's own pseudo-labels, the rest 0."""
        imgs, dt = self._images(images)
        n = imgs.shape[0]
        norm = self._norm(norm, n)
        if pad_to < self.K * 32:
            raise ValueError(f"pad_to must be >= K*32 = {self.K * 32}")
        out = torch.empty((n, pad_to), dtype=torch.float32, device=self.device)
        st = self._stream()
        with torch.cuda.device(self.device):
            for i in range(0, n, self.max_samples):
                m = min(self.max_samples, n - i)
                self._check(self.lib.b200al_badge_embed(self.ctx, self._ptr(imgs[i:i + m]), dt,
                                                        self._ptr(norm[i:i + m] if norm is not None else None), m,
                                                        float(dice_smooth), int(bool(dice_do_bg)),
                                                        self._ptr(out[i:i + m]), pad_to, st))
        return out

    def uncertainty_from_logits(self, logits: torch.Tensor, mode: str = "entropy", smooth: float = 1e-8):
        """logits [n,K,H,W] or [n,T,K,H,W] fp32 on device -> scores [n]."""
        if logits.dim() == 4:
            logits = logits.unsqueeze(1)
        n, T, K = logits.shape[:3]
        hw = logits.shape[3] * logits.shape[4]
        logits = logits.to(torch.float32).contiguous()
        out = torch.empty(n, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_uncertainty_from_logits(self.ctx, self._ptr(logits), n, T, K, hw,
                                                                _lib.MODES[mode], float(smooth), self._ptr(out),
                                                                self._stream()))
        return out

    def _reserve(self, nbytes: int) -> None:
        """Grow the context's scratch workspace (the only allocation of the rank / fused k-center paths; the C calls
        themselves never allocate)."""
        if nbytes > getattr(self, "_ws_bytes", 0):
            with torch.cuda.device(self.device):
                self._check(self.lib.b200al_workspace_reserve(self.ctx, int(nbytes)))
            self._ws_bytes = int(nbytes)

    # ------------------------------------------------------------ rank & cut
    def topk(self, scores: torch.Tensor, k: int, base_pos: int = 0):
        """k best (descending score, lower position on ties): (scores[k'], pos[k'] int64)."""
        scores = scores.to(device=self.device, dtype=torch.float32).contiguous()
        n = scores.numel()
        k = min(int(k), n)
        vals = torch.empty(k, dtype=torch.float32, device=self.device)
        pos = torch.empty(k, dtype=torch.int64, device=self.device)
        if k:
            self._reserve(self.lib.b200al_workspace_bytes_topk(n))
            with torch.cuda.device(self.device):
                self._check(self.lib.b200al_topk(self.ctx, self._ptr(scores), n, k, int(base_pos), self._ptr(vals),
                                                 self._ptr(pos), self._stream()))
        return vals, pos

    # --------------------------------------------------------------- k-center
    def kcenter(self, feats: torch.Tensor, n_init: int, budget: int, metric: str = "l2", return_vals: bool = False,
                criteria: str = "min"):
        """k-center greedy over rows of `feats` [n,D] (fp32, device); rows
        [0,n_init) are the initial centres (labeled set).  criteria "min"
        (farthest point from its nearest centre) or "mean" (largest mean
        distance to all centres), coreset_selector.py:33-36.  Returns row
        indices (int64, selection order)."""
        feats = feats.to(device=self.device, dtype=torch.float32).contiguous()
        n, D = feats.shape
        if n_init < 1:
            raise ValueError("k-center needs at least one labeled row (the reference falls back to random selection)")
        if criteria not in _lib.CRITERIA:
            raise RuntimeError(f"coreset_criteria {criteria} is undefined")  # coreset_selector.py:38-40
        budget = min(int(budget), n - n_init)
        picks = torch.empty(budget, dtype=torch.int64, device=self.device)
        vals = torch.empty(budget, dtype=torch.float32, device=self.device)
        if budget <= 0:
            return (picks, vals) if return_vals else picks
        st = self._stream()
        m, c = _lib.METRICS[metric], _lib.CRITERIA[criteria]
        with torch.cuda.device(self.device):
            if metric == "cosine":
                normed = torch.empty_like(feats)
                self._check(self.lib.b200al_normalize_rows(self.ctx, self._ptr(feats), self._ptr(normed), n, D, st))
                feats = normed
            # running minimum (fp32) or running sum (fp64) of the distances to the centres
            state = torch.empty(n, dtype=torch.float64 if criteria == "mean" else torch.float32, device=self.device)
            taken = torch.zeros(n, dtype=torch.uint8, device=self.device)
            taken[:n_init] = 1
            self._check(self.lib.b200al_kcenter_init(self.ctx, self._ptr(feats), n, D, self._ptr(feats), n_init, m, c,
                                                     self._ptr(state), 0, st))
            self._check(self.lib.b200al_kcenter_run(self.ctx, self._ptr(feats), n, D, m, c, n_init, self._ptr(state),
                                                    self._ptr(taken), budget, self._ptr(picks), self._ptr(vals), st))
        return (picks, vals) if return_vals else picks

    def normalize_rows(self, feats: torch.Tensor) -> torch.Tensor:
        """sklearn normalize() as cosine distances apply it (fp32, zero rows untouched)."""
        feats = feats.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty_like(feats)
        if feats.shape[0]:
            with torch.cuda.device(self.device):
                self._check(self.lib.b200al_normalize_rows(self.ctx, self._ptr(feats), self._ptr(out), feats.shape[0],
                                                           feats.shape[1], self._stream()))
        return out

    def kcenter_sharded(self, shard: torch.Tensor, centres: torch.Tensor, row_base: int, row_bases: list, budget: int,
                        metric: str, criteria: str, rank: int, world: int, epoch: int, mail_ptrs: list, row_ptrs: list,
                        timeout_ms: int = 20000) -> torch.Tensor:
        """k-center greedy with the rows sharded over `world` ranks (b200al_kcenter_run_sharded): `shard` = this
        rank's rows [n_local, D] (already normalised for cosine) living in peer-accessible memory, `centres` = the
        initial centres (labeled rows, the same on every rank).  Returns GLOBAL row indices, the same on every rank."""
        n_local, D = shard.shape
        centres = centres.to(device=self.device, dtype=torch.float32).contiguous()
        if criteria not in _lib.CRITERIA:
            raise RuntimeError(f"coreset_criteria {criteria} is undefined")
        m, c = _lib.METRICS[metric], _lib.CRITERIA[criteria]
        state = torch.empty(max(n_local, 1), dtype=torch.float64 if criteria == "mean" else torch.float32,
                            device=self.device)
        taken = torch.zeros(max(n_local, 1), dtype=torch.uint8, device=self.device)
        picks = torch.empty(budget, dtype=torch.int64, device=self.device)
        st = self._stream()
        PtrArr, I64Arr = C.c_void_p * world, C.c_int64 * (world + 1)
        with torch.cuda.device(self.device):
            if n_local:
                self._check(self.lib.b200al_kcenter_init(self.ctx, self._ptr(shard), n_local, D, self._ptr(centres),
                                                         centres.shape[0], m, c, self._ptr(state), 0, st))
            self._check(self.lib.b200al_kcenter_run_sharded(
                self.ctx, self._ptr(shard) if n_local else None, n_local, int(row_base), D, m, c, centres.shape[0],
                self._ptr(state), self._ptr(taken), int(budget), self._ptr(picks), None, world, rank, int(epoch),
                C.c_void_p(mail_ptrs[rank]), PtrArr(*mail_ptrs), PtrArr(*row_ptrs), I64Arr(*row_bases),
                int(timeout_ms), st))
        return picks

    def kcenter_fused(self, feats: torch.Tensor, loaded: torch.Tensor, loaded_weight: float, n_init: int, budget: int,
                      metric: str = "l2", criteria: str = "min", return_vals: bool = False):
        """k-center greedy on `w * D_loaded / sum(D_loaded) + (1 - w) * D_feats / sum(D_feats)` - CoresetSelector
        with externally loaded features and coreset_fusion="add" (coreset_selector.py:138-157).  Rows [0, n_init) of
        both matrices are the labeled set.  Row indices (int64, selection order)."""
        feats = feats.to(device=self.device, dtype=torch.float32).contiguous()
        loaded = loaded.to(device=self.device, dtype=torch.float32).contiguous()
        n = feats.shape[0]
        if loaded.shape[0] != n:
            raise ValueError("feats and loaded must have the same number of rows")
        if n_init < 1:
            raise ValueError("k-center needs at least one labeled row")
        if criteria not in _lib.CRITERIA:
            raise RuntimeError(f"coreset_criteria {criteria} is undefined")
        budget = min(int(budget), n - n_init)
        picks = torch.empty(budget, dtype=torch.int64, device=self.device)
        vals = torch.empty(budget, dtype=torch.float32, device=self.device)
        if budget > 0:
            self._reserve(self.lib.b200al_workspace_bytes_kcenter_fused(n, feats.shape[1], loaded.shape[1],
                                                                        _lib.METRICS[metric]))
            with torch.cuda.device(self.device):
                self._check(self.lib.b200al_kcenter_fused(
                    self.ctx, self._ptr(feats), feats.shape[1], self._ptr(loaded), loaded.shape[1], n, int(n_init),
                    _lib.METRICS[metric], _lib.CRITERIA[criteria], float(loaded_weight), float(1 - loaded_weight),
                    budget, self._ptr(picks), self._ptr(vals), self._stream()))
        return (picks, vals) if return_vals else picks

    def centre_distances(self, feats: torch.Tensor, centres: torch.Tensor, metric: str = "l2",
                         criteria: str = "min") -> torch.Tensor:
        """Per row of `feats` [n,D]: min (or mean) over the rows of `centres` [L,D] of the sklearn-fp32 distance -
        `pairwise_distances(feats, centres, metric).min(axis=1)` / `.mean(axis=1)` (kmean_selector.py:143-145,
        175-178) without the n x L matrix.  fp32 [n] on device."""
        feats = feats.to(device=self.device, dtype=torch.float32).contiguous()
        centres = centres.to(device=self.device, dtype=torch.float32).contiguous()
        n, D = feats.shape
        L = centres.shape[0]
        if criteria not in _lib.CRITERIA:
            raise RuntimeError(f"coreset_criteria {criteria} is undefined")
        st = self._stream()
        with torch.cuda.device(self.device):
            if metric == "cosine":
                fn, cn = torch.empty_like(feats), torch.empty_like(centres)
                self._check(self.lib.b200al_normalize_rows(self.ctx, self._ptr(feats), self._ptr(fn), n, D, st))
                self._check(self.lib.b200al_normalize_rows(self.ctx, self._ptr(centres), self._ptr(cn), L, D, st))
                feats, centres = fn, cn
            state = torch.empty(n, dtype=torch.float64 if criteria == "mean" else torch.float32, device=self.device)
            self._check(self.lib.b200al_kcenter_init(self.ctx, self._ptr(feats), n, D, self._ptr(centres), L,
                                                     _lib.METRICS[metric], _lib.CRITERIA[criteria], self._ptr(state),
                                                     0, st))
        return (state / L).to(torch.float32) if criteria == "mean" else state

    def standardize_rows(self, feats: torch.Tensor) -> torch.Tensor:
        """(x - mean) / std per row, population std (kmean_selector.py:95-97)."""
        feats = feats.to(device=self.device, dtype=torch.float32).contiguous()
        out = torch.empty_like(feats)
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_standardize_rows(self.ctx, self._ptr(feats), self._ptr(out), feats.shape[0],
                                                         feats.shape[1], self._stream()))
        return out

    def kmeans_pp(self, X: torch.Tensor, weights: torch.Tensor, first_centre: int, uniforms: torch.Tensor,
                  n_clusters: int, n_local_trials: int) -> torch.Tensor:
        """scikit-learn's weighted k-means++ seeding on rows of X [n,D] (fp32, device) with the host-drawn random
        numbers (`first_centre`, `uniforms` [(n_clusters-1), n_local_trials] float64).  Row indices, pick order."""
        X = X.to(device=self.device, dtype=torch.float32).contiguous()
        n, D = X.shape
        weights = weights.to(device=self.device, dtype=torch.float32).contiguous()
        uniforms = uniforms.to(device=self.device, dtype=torch.float64).contiguous()
        if weights.shape != (n,) or uniforms.numel() != max(0, n_clusters - 1) * n_local_trials:
            raise ValueError("weights must be [n] and uniforms [(n_clusters-1) * n_local_trials]")
        picks = torch.empty(n_clusters, dtype=torch.int64, device=self.device)
        nbytes = self.lib.b200al_kmeans_pp_scratch_bytes(n, n_local_trials)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_kmeans_pp(self.ctx, self._ptr(X), n, D, self._ptr(weights), int(first_centre),
                                                  self._ptr(uniforms), int(n_clusters), int(n_local_trials),
                                                  self._ptr(scratch), self._ptr(picks), self._stream()))
        return picks

    # ---------------------------------------------------------- introspection
    def kernel_launches(self) -> int:
        c = C.c_int64()
        self._check(self.lib.b200al_kernel_launches(self.ctx, C.byref(c)))
        return c.value

    def set_timing(self, enable: bool):
        self._check(self.lib.b200al_set_timing(self.ctx, int(enable)))

    def conv_time_ms(self):
        """(summed tcgen05 conv kernel ms, timed launches, their FLOPs) since set_timing(True)."""
        ms, n, fl = C.c_float(), C.c_int64(), C.c_double()
        self._check(self.lib.b200al_conv_time_ms(self.ctx, C.byref(ms), C.byref(n), C.byref(fl)))
        return ms.value, n.value, fl.value

    def layer_times_ms(self):
        arr = (C.c_float * 21)()
        self._check(self.lib.b200al_debug_layer_times(self.ctx, arr, 21))
        return list(arr)

    # -------------------------------------------------------------- test hook
    def debug_layer_info(self, i: int):
        arr = (C.c_int * 12)()
        self._check(self.lib.b200al_debug_layer_info(self.ctx, i, arr))
        keys = ("kind", "cin", "cout", "stride", "inH", "inW", "in_pitch", "in_c_off", "outH", "outW", "out_pitch",
                "out_c_off")
        return dict(zip(keys, list(arr)))

    def debug_layer(self, i: int, impl: int, x: torch.Tensor, out: torch.Tensor, mask: torch.Tensor | None = None):
        with torch.cuda.device(self.device):
            self._check(self.lib.b200al_debug_layer(self.ctx, i, impl, self._ptr(x), self._ptr(out), x.shape[0],
                                                    self._ptr(mask), self._stream()))


# FLOPs per sample of the tensor-core layers (2*MAC; SURVEY.md 8a'), H=W=256
def conv_flops_per_sample(H: int = 256, W: int = 256, encoder_only: bool = False) -> float:
    f = 0.0
    px = H * W
    ch = CHANNELS
    f += 2.0 * px * 32 * 32 * 9                      # enc0.1
    for l in range(1, 5):
        px_l = px // (4 ** l)
        f += 2.0 * px_l * ch[l] * ch[l - 1] * 9      # stride-2 conv
        f += 2.0 * px_l * ch[l] * ch[l] * 9
    if encoder_only:
        return f
    for l in range(4):
        cin, cout = ch[4 - l], ch[3 - l]
        px_in = px // (4 ** (4 - l))
        f += 2.0 * px_in * 4 * cout * cin            # transposed conv
        px_out = px_in * 4
        f += 2.0 * px_out * cout * (2 * cout) * 9
        f += 2.0 * px_out * cout * cout * 9
    return f
