"""Drop-in selection strategies with the reference's plugin interface.

Mirrors `ActiveSelector.select_next_batch(active_dataset, select_num, model,
device) -> list[case_name]` (src/activelearning/active_selector.py:10-19) and
the constructor kwargs `ALTrainer._setup_active_selector` passes
(src/training/al_trainer.py:802-844).  Scores, embeddings, rank & cut and
k-center all run in libb200al on the GPU; Python only moves pool chunks to the
device and maps positions back to case names.

Objects consumed (duck-typed, reference types or mirrors): `active_dataset`
with get_size / get_pool_dataset / get_train_dataset; datasets with
`.image_idx: list[str]` and `__getitem__ -> {"image": float32 [1,H,W], ...}`.
Fast paths: (1) if the wrapped base dataset exposes `images_u8` ([N,H,W] uint8
tensor, CPU or CUDA) and `names`, chunks are sliced from it directly; (2) for
datasets whose raw pixels can be read (pool_cache.py: the reference's FUGC /
BUSI datasets, or anything with `b200_raw_u8`) the raw bytes are cached in HBM
once per run and resize + z-score run on the GPU.  Otherwise a DataLoader with
the reference's settings feeds fp32 batches.

Every `select_next_batch` leaves `self.last_timing` (seconds, pool size,
images/s) and logs it on the "b200al" logger - the reference's trainer times
epochs and iterations but not acquisition (al_trainer.py:1148-1202).
"""
from __future__ import annotations

import functools
import logging
import time
from abc import ABC, abstractmethod
from typing import Any, Iterator

import torch

from . import dist as D
from .engine import Engine, validate_state_dict
from .pool_cache import pool_cache_for

log = logging.getLogger("b200al")


class ActiveSelector(ABC):
    """Same abstract interface as the reference's ActiveSelector."""

    @abstractmethod
    def select_next_batch(self, active_dataset, select_num: int, model, device) -> list[Any]:
        ...


_ENGINES: dict = {}


def get_engine(device, num_out: int, H: int, W: int, precision: str, max_samples: int) -> Engine:
    device = torch.device(device)
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, num_out, H, W, precision, max_samples)
    eng = _ENGINES.get(key)
    if eng is None:
        eng = _ENGINES[key] = Engine(torch.device("cuda", idx), num_out, H, W, max_samples, precision)
    return eng


def _cold_start(active_dataset, select_num: int) -> list:
    """labeled_size == 0 branch, verbatim call sequence so the seeded global
    CPU RNG stream matches the reference (entropy_selector.py:64-71)."""
    _, pool_size = active_dataset.get_size()
    scores = torch.rand(pool_size)
    _, indices = torch.sort(scores, descending=True)
    indices = D.broadcast_cpu(indices[:select_num])  # multi-GPU: rank 0's draw decides (rank RNG states may differ)
    return [active_dataset.pool_dataset.image_idx[id] for id in indices]


class PoolSource:
    """Feeds device-resident image chunks of a dataset in position order."""

    def __init__(self, dataset, device, batch_size: int, num_workers: int, pin_memory: bool, pool_cache: str = "auto"):
        self.ds, self.device = dataset, torch.device(device)
        self.batch_size, self.num_workers, self.pin_memory = batch_size, num_workers, pin_memory
        base = getattr(dataset, "dataset", None)
        self.resident = None
        self.cache = None
        if base is not None and hasattr(dataset, "case_name_to_idx"):
            if hasattr(base, "images_u8"):
                self.resident = base.images_u8
            else:
                self.cache = pool_cache_for(base, self.device, pool_cache)

    def image_hw(self) -> tuple[int, int]:
        if self.resident is not None:
            return tuple(self.resident.shape[-2:])
        if self.cache is not None:
            return self.cache.out_hw
        if len(self.ds) == 0:
            return (256, 256)
        return tuple(self.ds[0]["image"].shape[-2:])

    def chunks(self, positions: list[int], chunk: int, eng: Engine | None = None
               ) -> Iterator[tuple[int, torch.Tensor, torch.Tensor | None]]:
        """Yields (offset into `positions`, images on device, per-image (mean, std) or None) with one chunk of
        host->device copy in flight ahead of the consumer."""
        if not positions:
            return
        if self.cache is not None:
            rows = torch.tensor([self.ds.case_name_to_idx[self.ds.image_idx[p]] for p in positions], dtype=torch.int64,
                                device=self.cache.device)
            for i in range(0, len(positions), chunk):
                imgs, norm = self.cache.gather(eng, rows[i:i + chunk])
                yield i, imgs, norm
            return
        if self.resident is not None:
            rows = torch.tensor([self.ds.case_name_to_idx[self.ds.image_idx[p]] for p in positions], dtype=torch.int64)
            if self.resident.is_cuda:
                rows = rows.to(self.resident.device)
                for i in range(0, len(positions), chunk):
                    yield i, self.resident.index_select(0, rows[i:i + chunk]).to(self.device), None
                return
            copy_stream = torch.cuda.Stream(self.device)
            cur = torch.cuda.current_stream(self.device)
            contiguous = bool((rows[1:] - rows[:-1] == 1).all()) if len(rows) > 1 else True

            def stage(i):
                if contiguous:
                    host = self.resident[int(rows[i]): int(rows[i]) + len(rows[i:i + chunk])]
                else:
                    host = self.resident.index_select(0, rows[i:i + chunk])
                if not host.is_pinned() and self.pin_memory:
                    host = host.pin_memory()
                with torch.cuda.stream(copy_stream):
                    dev = host.to(self.device, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(copy_stream)
                return dev, ev, host

            nxt = stage(0)
            for i in range(0, len(positions), chunk):
                dev, ev, _host = nxt
                if i + chunk < len(positions):
                    nxt = stage(i + chunk)
                cur.wait_event(ev)
                dev.record_stream(cur)
                yield i, dev, None
            return
        # generic path: the reference's own DataLoader settings (entropy_selector.py:30-35)
        from torch.utils.data import DataLoader, Subset
        loader = DataLoader(Subset(self.ds, positions), batch_size=min(chunk, max(1, self.batch_size)),
                            num_workers=self.num_workers, pin_memory=self.pin_memory)
        off = 0
        for batch in loader:
            img = batch["image"].to(self.device, non_blocking=True).to(torch.float32)
            yield off, img, None
            off += img.shape[0]


class _Stages:
    """Per-stage wall time of one select_next_batch (weights / score / embed / gather / rank / kcenter ...).  Off by
    default: with `profile_stages=True` the device is synchronised at every stage boundary, which costs overlap."""

    def __init__(self, enabled: bool, device):
        self.enabled, self.times = enabled, {}
        self.device = torch.device(device)
        self._name, self._t0 = None, 0.0

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def start(self, name: str):
        if not self.enabled:
            return
        self.stop()
        self._sync()
        self._name, self._t0 = name, time.perf_counter()

    def stop(self):
        if self.enabled and self._name is not None:
            self._sync()
            self.times[self._name] = self.times.get(self._name, 0.0) + time.perf_counter() - self._t0
            self._name = None


def _timed(fn):
    @functools.wraps(fn)
    def wrapper(self, active_dataset, select_num, model, device):
        dev = torch.device(device)
        self._stages = _Stages(self.profile_stages, dev)
        t0 = time.perf_counter()
        out = fn(self, active_dataset, select_num, model, device)
        self._stages.stop()
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        labeled, pool = active_dataset.get_size()
        self.last_timing = {"seconds": dt, "pool": pool, "labeled": labeled, "selected": len(out),
                            "images_per_s": pool / dt if dt > 0 else float("inf"),
                            "stages": dict(self._stages.times)}
        log.info("%s: %d of %d pool images selected in %.3f s (%.0f images/s, %d labeled)", type(self).__name__,
                 len(out), pool, dt, self.last_timing["images_per_s"], labeled)
        return out

    return wrapper


class _B200Base(ActiveSelector):
    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8, *,
                 precision: str = "mixed", chunk_samples: int = 64, pool_cache: str = "auto",
                 sync_weights: bool = True, profile_stages: bool = False):
        self.batch_size, self.num_workers, self.pin_memory = batch_size, num_workers, pin_memory
        self.smooth = smooth
        self.precision = precision
        self.chunk_samples = chunk_samples
        if pool_cache not in ("auto", "off"):
            raise ValueError("pool_cache must be 'auto' or 'off'")
        self.pool_cache = pool_cache
        self.sync_weights = bool(sync_weights)  # multi-GPU: every rank scores with rank 0's weights
        self.last_timing: dict | None = None
        self.profile_stages = bool(profile_stages)
        self._stages = _Stages(False, "cpu")

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        if "select_next_batch" in cls.__dict__:  # acquisition-time logging for every concrete selector
            cls.select_next_batch = _timed(cls.__dict__["select_next_batch"])

    def _source(self, dataset, device) -> "PoolSource":
        return PoolSource(dataset, device, self.batch_size, self.num_workers, self.pin_memory, self.pool_cache)

    def _engine(self, model, device, hw) -> Engine:
        self._stages.start("weights")
        sd = model.state_dict()
        K = validate_state_dict(sd)
        eng = get_engine(device, K, hw[0], hw[1], self.precision, self.chunk_samples)
        model.eval()  # the reference's own side effect (entropy_selector.py:42)
        # every call: the trainer rebuilds / reloads the model each round (al_trainer.py:1048-1083)
        if self.sync_weights and D.world()[1] > 1:
            eng.load_model(model, D.broadcast_state_dict(sd, eng.device))
        else:
            eng.load_model(model)
        return eng


class _B200Uncertainty(_B200Base):
    mode = "entropy"

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8, *,
                 mc_repeats: int = 1, dropout_prob: float | None = None, mc_seed: int = 0, **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)
        self.mc_repeats, self.dropout_prob, self.mc_seed = int(mc_repeats), dropout_prob, int(mc_seed)

    def _score_positions(self, eng: Engine, src: PoolSource, positions: list[int]):
        T = self.mc_repeats
        p = 0.0
        if T > 1:
            p = self.dropout_prob if self.dropout_prob is not None else (eng.dropout_p or 0.0)
        out = torch.empty(len(positions), dtype=torch.float32, device=eng.device)
        per = max(1, eng.max_samples // T)
        for off, imgs, norm in src.chunks(positions, per, eng):
            # MC masks are keyed by pool position so any sharding gives the same scores
            out[off:off + imgs.shape[0]] = eng.scores(imgs, self.mode, T=T, drop_p=p, seed=self.mc_seed,
                                                      base_pos=positions[off], smooth=self.smooth, norm=norm)
        return out

    def _scores_local(self, active_dataset, model, device):
        pool = active_dataset.get_pool_dataset()
        rank, ws = D.world()
        lo, hi = D.shard_range(len(pool), rank, ws)
        src = self._source(pool, device)
        eng = self._engine(model, device, src.image_hw())
        self._stages.start("score")
        return eng, self._score_positions(eng, src, list(range(lo, hi))), lo, hi

    def cal_scores(self, active_dataset, model, device):
        """(scores [device tensor], case names) of this rank's pool shard, in
        pool order - the whole pool when not distributed, like the reference's
        cal_scores (entropy_selector.py:24-54)."""
        _eng, scores, lo, hi = self._scores_local(active_dataset, model, device)
        return scores, active_dataset.get_pool_dataset().image_idx[lo:hi]

    def select_next_batch(self, active_dataset, select_num: int, model, device):
        labeled_size, pool_size = active_dataset.get_size()
        if labeled_size == 0:
            return _cold_start(active_dataset, select_num)
        eng, scores, lo, _hi = self._scores_local(active_dataset, model, device)
        self._stages.start("rank")
        vals, pos = eng.topk(scores, select_num, base_pos=lo)
        if D.world()[1] > 1:
            vals, pos = D.global_topk(vals, pos, select_num)
        names = active_dataset.get_pool_dataset().image_idx
        return [names[p] for p in pos.tolist()]


class B200EntropySelector(_B200Uncertainty):
    """EntropySelector (src/activelearning/entropy_selector.py:11)."""
    mode = "entropy"


class B200MarginSelector(_B200Uncertainty):
    """MarginSelector (src/activelearning/margin_selector.py:11)."""
    mode = "margin"


class B200ConfidenceSelector(_B200Uncertainty):
    """ConfidenceSelector (src/activelearning/confidence_selector.py:11)."""
    mode = "confidence"


class B200DiceUncertaintySelector(_B200Uncertainty):
    """Dice-uncertainty (BASELINE.json north star, kernel 3; not a reference selector): score =
    DiceLoss(p, one_hot(argmax p)) of the (T-averaged) softmax map against its own hard prediction, with the
    reference's DiceLoss semantics (src/losses/dice_loss.py:32-72: do_bg=True, batch=False, squared=False,
    smooth 1e-5 as al_trainer.py:785-791 builds it).  `smooth` is DiceLoss.smooth here."""
    mode = "dice"

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-5, **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)


class B200MCDropoutEntropySelector(_B200Uncertainty):
    """Entropy of the mean of T MC-dropout softmax maps (not in the reference;
    composed from its UNet + Dropout2d + entropy formula)."""
    mode = "entropy"

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8, *,
                 mc_repeats: int = 8, **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, mc_repeats=mc_repeats, **kw)


def _embed_positions(eng: Engine, src: PoolSource, positions: list[int]) -> torch.Tensor:
    out = torch.empty((len(positions), 512), dtype=torch.float32, device=eng.device)
    for off, imgs, norm in src.chunks(positions, eng.max_samples, eng):
        out[off:off + imgs.shape[0]] = eng.embed(imgs, norm=norm)
    return out


KERNEL_WIDTHS = (128, 256, 512, 1024, 1536, 2048)  # row widths the k-center / k-means++ kernels are built for


def _pad_to_kernel_width(x: torch.Tensor) -> torch.Tensor:
    """Zero-pad the columns of x to the next supported width (zero columns change neither l2 nor cosine distances)."""
    width = next((w for w in KERNEL_WIDTHS if w >= x.shape[1]), None)
    if width is None:
        raise NotImplementedError(f"feature rows of dimension {x.shape[1]} (> {KERNEL_WIDTHS[-1]}) are not supported")
    return torch.nn.functional.pad(x, (0, width - x.shape[1])) if width != x.shape[1] else x


def _read_feature_file(feature_path, case) -> "np.ndarray":
    """The `feature` dataset of `{feature_path}/{case}.h5` (coreset_selector.py:110-119, kmean_selector.py:75-84;
    needs h5py like the reference), or - where h5py is not installed - `{case}.npy` / `{case}.npz` (key "feature")
    holding the same array."""
    import numpy as np
    from pathlib import Path
    base = Path(feature_path)
    h5 = base / f"{case}.h5"
    if h5.exists():
        try:
            import h5py
        except ImportError as e:
            raise NotImplementedError(f"{h5} needs h5py (not installed); store the row as {case}.npy / {case}.npz "
                                      "or pass feature_dict") from e
        with h5py.File(h5, "r") as f:
            return np.asarray(f["feature"][:])
    npy = base / f"{case}.npy"
    if npy.exists():
        return np.load(npy)
    npz = base / f"{case}.npz"
    if npz.exists():
        with np.load(npz) as z:
            return np.asarray(z["feature"])
    raise FileNotFoundError(f"no feature file for case {case!r} under {base} (.h5 / .npy / .npz)")


def _loaded_feature_rows(names, feature_dict=None, feature_path=None):
    """[len(names), D2] float32 array of externally loaded (foundation-model) features: rows of `feature_dict`, or
    one file per case under `feature_path` (`_read_feature_file`)."""
    import numpy as np
    if feature_dict is not None:
        return np.stack([np.asarray(feature_dict[n], dtype=np.float32).reshape(-1) for n in names], axis=0)
    return np.stack([np.asarray(_read_feature_file(feature_path, n), dtype=np.float32).reshape(-1) for n in names], axis=0)


def _cat_scale(d1: int, d2: int, weight: float) -> "np.float32":
    """`np.sqrt(D1 / D2 * loaded_feature_weight)` (coreset_selector.py:166-170, kmean_selector.py:113-116) as the
    float32 the reference multiplies with: it pins numpy < 2 (pyproject.toml:21), where `float32_array *
    np.float64_scalar` stays float32 and the scalar is cast to float32 first (value-based casting)."""
    import numpy as np
    return np.float32(np.sqrt(d1 / d2 * weight))


def _require_finite(t: torch.Tensor, what: str) -> None:
    """The selection kernels compare with `>` / `<`, which NaN never satisfies (a diverged model, or a constant row
    through standardize_rows' division by std = 0): refuse instead of returning an arbitrary selection."""
    if not bool(torch.isfinite(t).all()):
        raise RuntimeError(f"non-finite values in the {what}: the selection would be arbitrary")


def _check_picks(picks: list, what: str) -> list:
    if any(i < 0 for i in picks):
        raise RuntimeError(f"{what}: no selectable row left (non-finite distances?)")
    return picks


class B200CoresetSelector(_B200Base):
    """CoresetSelector (src/activelearning/coreset_selector.py:54) for metric in {l2, cosine}, coreset_criteria in
    {min, mean}; model features alone, or fused with externally loaded features by coreset_fusion="add"
    (:138-157: each distance matrix divided by its global sum, mixed with `loaded_feature_weight`).  Loaded features
    come from `feature_path` (`{case}.h5`, needs h5py like the reference) or from `feature_dict` (case name -> row;
    an extension for callers that already hold them in memory, like KMeanSelector's).  coreset_fusion="cat" (:158-175)
    concatenates the embeddings with the loaded rows scaled by sqrt(D1 / D2 * loaded_feature_weight) - in float32, the
    reference's behaviour under its pinned numpy < 2 (`_cat_scale`) - and runs k-center on the joint rows."""

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8,
                 metric: str = "cosine", coreset_criteria: str = "min", coreset_fusion: str = "add",
                 feature_path=None, loaded_feature_weight: float = 0.0, feature_dict=None, shard_kcenter: bool = True,
                 **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)
        self.shard_kcenter = bool(shard_kcenter)  # multi-GPU: sharded greedy loop (False: gather rows, run replicated)
        if metric not in ("l2", "cosine"):
            raise NotImplementedError(f"metric {metric!r}: only 'l2' and 'cosine' are supported")
        if coreset_criteria not in ("min", "mean"):
            raise RuntimeError(f"coreset_criteria {coreset_criteria} is undefined")  # coreset_selector.py:38-40
        self.has_loaded = feature_path is not None or feature_dict is not None
        if coreset_fusion not in ("add", "cat"):
            raise NotImplementedError(f"coreset_fusion {coreset_fusion!r}: only 'add' and 'cat' exist in the reference")
        if feature_path is not None and feature_dict is None:
            from pathlib import Path
            feature_path = Path(feature_path)
        self.metric = metric
        self.coreset_criteria = coreset_criteria
        self.coreset_fusion = coreset_fusion
        self.feature_path, self.feature_dict = feature_path, feature_dict
        self.loaded_feature_weight = float(loaded_feature_weight)

    def _kcenter_sharded(self, eng: Engine, emb_l, emb_local, labeled_size, pool_size, select_num, rank, ws) -> list:
        """Multi-GPU k-center without gathering the embeddings: every rank keeps the running state of its own pool
        rows, and one kernel per rank runs all greedy steps, exchanging per step only its best (value, row) candidate
        and reading the chosen centre from its owner's shard over NVLink (csrc/kcenter.cuh, kcenter_persist_kernel).
        Global rows: 0..L-1 labeled (coreset_selector.py:85-88), L + p = pool position p."""
        self._stages.start("kcenter")
        # collective check: a NaN in ONE shard must make every rank raise (not leave the others waiting at the barrier)
        ok = torch.tensor([float(bool(torch.isfinite(emb_l).all()) and bool(torch.isfinite(emb_local).all()))],
                          device=eng.device)
        torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)
        if not bool(ok.item()):
            raise RuntimeError("non-finite values in the embeddings: the selection would be arbitrary")
        if self.metric == "cosine":
            emb_l, emb_local = eng.normalize_rows(emb_l), eng.normalize_rows(emb_local)
        bases = [labeled_size + D.shard_range(pool_size, r, ws)[0] for r in range(ws)] + [labeled_size + pool_size]
        cap = max(bases[r + 1] - bases[r] for r in range(ws))
        pb = D.peer_buffers(eng.device, cap, emb_local.shape[1], eng.lib.b200al_kcenter_mailbox_bytes(ws))
        shard = pb.rows(emb_local.shape[0])
        shard.copy_(emb_local)
        epoch = pb.next_epoch()
        torch.cuda.synchronize(eng.device)
        torch.distributed.barrier()   # every shard is in place before any rank's kernel reads a peer's rows
        budget = min(int(select_num), pool_size)
        picks = eng.kcenter_sharded(shard, emb_l, bases[rank], bases, budget, self.metric, self.coreset_criteria,
                                    rank, ws, epoch, pb.mail_ptrs(), pb.row_ptrs())
        return picks.tolist()

    def select_next_batch(self, active_dataset, select_num: int, model, device):
        import numpy as np
        labeled_size, pool_size = active_dataset.get_size()
        labeled, pool = active_dataset.get_train_dataset(), active_dataset.get_pool_dataset()
        if labeled_size == 0:
            if self.loaded_feature_weight == 0 or not self.has_loaded:
                return _cold_start(active_dataset, select_num)  # coreset_selector.py:193-200, 211-217
            # :201-210 - k-means++ seeding over the loaded features alone, numpy's global RNG
            if select_num <= 0:
                return []
            if pool_size < select_num:
                raise ValueError(f"n_samples={pool_size} should be >= n_clusters={select_num}.")
            eng = get_engine(device, 2, 256, 256, self.precision, 1)
            rows = torch.from_numpy(_loaded_feature_rows(pool.image_idx, self.feature_dict, self.feature_path))
            X = _pad_to_kernel_width(rows).to(eng.device)
            _require_finite(X, "loaded features")
            rs = np.random.mtrand._rand
            trials = 2 + int(np.log(select_num))
            w = np.ones(pool_size, dtype=np.float32)
            first = int(rs.choice(pool_size, p=w / w.sum()))
            uniforms = np.stack([rs.uniform(size=trials) for _ in range(select_num - 1)]) if select_num > 1 \
                else np.zeros((0, trials))
            first, uniforms = D.broadcast_draws(first, uniforms, eng.device)
            picks = eng.kmeans_pp(X, torch.from_numpy(w), first, torch.from_numpy(uniforms), select_num, trials)
            return [pool.image_idx[i] for i in picks.tolist()]
        all_list = list(labeled.image_idx) + list(pool.image_idx)  # labeled rows first (coreset_selector.py:85-88)
        rank, ws = D.world()
        lo, hi = D.shard_range(pool_size, rank, ws)
        psrc = self._source(pool, device)
        lsrc = self._source(labeled, device)
        eng = self._engine(model, device, psrc.image_hw())
        self._stages.start("embed")
        emb_l = _embed_positions(eng, lsrc, list(range(labeled_size)))
        emb_local = _embed_positions(eng, psrc, list(range(lo, hi)))
        if ws > 1 and not self.has_loaded and self.shard_kcenter:
            picks = self._kcenter_sharded(eng, emb_l, emb_local, labeled_size, pool_size, select_num, rank, ws)
            return [all_list[i] for i in _check_picks(picks, "k-center")]
        self._stages.start("gather")
        emb_p = D.gather_rows(emb_local)
        feats = torch.cat([emb_l, emb_p], 0)
        _require_finite(feats, "embeddings")
        self._stages.start("kcenter")
        if self.has_loaded:
            loaded = torch.from_numpy(_loaded_feature_rows(all_list, self.feature_dict, self.feature_path))
            _require_finite(loaded, "loaded features")
            if self.coreset_fusion == "add":
                picks = eng.kcenter_fused(feats, loaded, self.loaded_feature_weight, labeled_size, select_num,
                                          self.metric, self.coreset_criteria)
            else:  # "cat" (:158-175): one distance matrix over [embeddings | scaled loaded rows], no /sum
                scale = _cat_scale(feats.shape[1], loaded.shape[1], self.loaded_feature_weight)
                joint = torch.cat([feats, (loaded * torch.tensor(scale)).to(eng.device)], 1)
                picks = eng.kcenter(_pad_to_kernel_width(joint), labeled_size, select_num, self.metric,
                                    criteria=self.coreset_criteria)
        else:
            picks = eng.kcenter(feats, labeled_size, select_num, self.metric, criteria=self.coreset_criteria)
        return [all_list[i] for i in _check_picks(picks.tolist(), "k-center")]


class B200KMeanSelector(_B200Base):
    """KMeanSelector (src/activelearning/kmean_selector.py:19-196): row-standardised `get_enc_feature` embeddings
    (:95-97) and / or externally loaded features (:75-90, 101-119), sample weight = min / mean distance of each pool
    row to the labeled rows (:143-145, 174-186) and scikit-learn's weighted k-means++ seeding over the pool (:190-192).

    The random numbers come from numpy's GLOBAL RandomState in sklearn's call order (one `choice`, then
    `uniform(size=2+int(log k))` per further centre), so a seeded run consumes the same stream as the reference.
    Loaded features: `feature_dict` (case name -> row) or `feature_path` (one `{case}.h5` per case like the reference,
    or `.npy` / `.npz` where h5py is missing).  With `loaded_feature_only=True` (the demo's model-free mode,
    entry/demo/serve.py:125-134) they replace the embeddings; otherwise the standardised loaded rows are scaled by
    sqrt(512 / D2 * loaded_feature_weight) - in float32, the reference's behaviour under its pinned numpy < 2
    (`_cat_scale`) - and appended to the standardised embeddings (:106-121).
    Embedding, standardisation, pool-to-labeled distances and the whole seeding loop run on the GPU; the O(n)
    weight normalisation and the draws stay on the host in the reference's own numpy / torch calls.  Names are
    returned in pick order, duplicates dropped (the reference returns `list(set(...))`, :194).  Without labeled data
    the weights are uniform (:187-188) - this selector has no random cold-start branch.  Picks are bit-pinned to
    scikit-learn 1.9.0's `_kmeans_plusplus` (fp32 sequential cumsum; <= 1.7 used a float64 `stable_cumsum`)."""

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8,
                 metric: str = "cosine", feature_path=None, feature_dict=None, coreset_criteria: str = "min",
                 loaded_feature_weight: float = 1.0, loaded_feature_only: bool = False, sharp_factor: float = 1.0,
                 softmax: bool = False, **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)
        if metric not in ("l2", "cosine"):
            raise NotImplementedError(f"metric {metric!r}: only 'l2' and 'cosine' are supported")
        if coreset_criteria not in ("min", "mean"):
            raise RuntimeError(f"coreset_criteria {coreset_criteria} is undefined")
        self.has_loaded = feature_path is not None or bool(feature_dict)  # `elif self.feature_dict:` (:86)
        if loaded_feature_only and not self.has_loaded:
            # the reference would reach np.concatenate([]) (:123) and raise there
            raise ValueError("loaded_feature_only=True needs feature_path or feature_dict")
        if feature_path is not None:
            from pathlib import Path
            feature_path = Path(feature_path)
        self.feature_path, self.feature_dict = feature_path, feature_dict
        self.loaded_feature_only = bool(loaded_feature_only)
        self.metric, self.coreset_criteria = metric, coreset_criteria
        self.sharp_factor, self.softmax = sharp_factor, bool(softmax)
        self.loaded_feature_weight = loaded_feature_weight

    def _loaded_rows(self, eng: Engine, names) -> torch.Tensor:
        """Loaded rows for `names`, standardised per row (:101-105); on device, unpadded."""
        # feature_path wins over feature_dict like the reference's if / elif (:75, :86)
        rows = _loaded_feature_rows(names, None if self.feature_path is not None else self.feature_dict,
                                    self.feature_path)
        return eng.standardize_rows(torch.from_numpy(rows).to(eng.device))

    def _features(self, eng: Engine, emb: torch.Tensor | None, names) -> torch.Tensor:
        """get_features (:48-127) after the embedding loop: [standardised embeddings | scaled standardised loaded
        rows], zero-padded to a kernel width."""
        parts = []
        if emb is not None:
            parts.append(eng.standardize_rows(emb))
        if self.has_loaded:
            loaded = self._loaded_rows(eng, names)
            if emb is not None:
                loaded = loaded * torch.tensor(_cat_scale(emb.shape[1], loaded.shape[1], self.loaded_feature_weight),
                                               device=eng.device)
            parts.append(loaded)
        x = parts[0] if len(parts) == 1 else torch.cat(parts, 1)
        _require_finite(x, "standardised features (a constant row has std = 0)")
        return _pad_to_kernel_width(x)

    def select_next_batch(self, active_dataset, select_num: int, model, device):
        import numpy as np
        labeled_size, pool_size = active_dataset.get_size()
        if select_num <= 0:
            return []
        if pool_size < select_num:  # sklearn's own check inside kmeans_plusplus
            raise ValueError(f"n_samples={pool_size} should be >= n_clusters={select_num}.")
        labeled, pool = active_dataset.get_train_dataset(), active_dataset.get_pool_dataset()
        rank, ws = D.world()
        if self.loaded_feature_only:
            eng = get_engine(device, 2, 256, 256, self.precision, 1)  # kernels only; no model involved
            X = self._features(eng, None, pool.image_idx)
            Xl = self._features(eng, None, labeled.image_idx) if labeled_size > 0 else None
        else:
            lo, hi = D.shard_range(pool_size, rank, ws)
            psrc = self._source(pool, device)
            eng = self._engine(model, device, psrc.image_hw())
            X = self._features(eng, D.gather_rows(_embed_positions(eng, psrc, list(range(lo, hi)))), pool.image_idx)
            Xl = None
            if labeled_size > 0:
                lsrc = self._source(labeled, device)
                Xl = self._features(eng, _embed_positions(eng, lsrc, list(range(labeled_size))), labeled.image_idx)
        if Xl is not None:
            w = eng.centre_distances(X, Xl, self.metric, self.coreset_criteria).cpu()
            if self.softmax:  # kmean_selector.py:180-183
                w = (w * self.sharp_factor).softmax(0).numpy()
            else:             # :184-186
                w = w.numpy() ** self.sharp_factor
                w = w / w.sum()
        else:
            w = np.ones(pool_size, dtype=np.float32)  # sklearn _check_sample_weight(None)
        w = np.asarray(w, dtype=np.float32)
        rs = np.random.mtrand._rand  # check_random_state(None): numpy's global RandomState
        trials = 2 + int(np.log(select_num))
        first = int(rs.choice(pool_size, p=w / w.sum()))
        uniforms = np.stack([rs.uniform(size=trials) for _ in range(select_num - 1)]) if select_num > 1 \
            else np.zeros((0, trials))
        # every rank consumed its own stream; rank 0's draws decide so that all ranks agree
        first, uniforms = D.broadcast_draws(first, uniforms, eng.device)
        picks = eng.kmeans_pp(X, torch.from_numpy(w), first, torch.from_numpy(uniforms), select_num, trials).tolist()
        seen, out = set(), []
        for i in picks:
            if i not in seen:
                seen.add(i)
                out.append(pool.image_idx[i])
        return out


def _dice_params(dice_loss) -> tuple[float, bool]:
    """(smooth, do_bg) of the trainer's DiceLoss instance (src/losses/dice_loss.py:7-30); anything the closed form in
    csrc/badge.cuh does not cover is rejected."""
    if dice_loss is None:
        return 1e-5, True  # al_trainer.py:785-791
    if not getattr(dice_loss, "softmax", True) or getattr(dice_loss, "squared", False):
        raise NotImplementedError("BADGE: only DiceLoss(softmax=True, squared=False) is supported")
    return float(getattr(dice_loss, "smooth", 1e-5)), bool(getattr(dice_loss, "do_bg", True))


def _check_ce(ce_loss) -> None:
    if ce_loss is None:
        return
    import torch.nn as nn
    if not isinstance(ce_loss, nn.CrossEntropyLoss) or ce_loss.weight is not None or ce_loss.reduction != "mean" \
            or getattr(ce_loss, "label_smoothing", 0.0) != 0.0:
        raise NotImplementedError("BADGE: only a plain torch.nn.CrossEntropyLoss() (mean reduction, no class weights, "
                                  "no label smoothing) is supported")


class B200BADGESelector(_B200Base):
    """BADGESelector (src/activelearning/badge_selector.py:37-128): per-image gradient embedding of
    `decoder.seg_output.weight` under CE + Dice against the model's own pseudo-labels, then scikit-learn's
    (unweighted) k-means++ seeding over the pool with numpy's global RNG.  The reference runs one autograd backward
    per image (batch_size is forced to 1, al_trainer.py:875); here the head-weight gradient is evaluated in closed
    form from the last feature map in two fused passes (csrc/badge.cuh), chunked like the other selectors.
    `dice_loss` / `ce_loss`: the trainer's loss modules (al_trainer.py:872-874) - only their hyper-parameters are read.
    """

    def __init__(self, dice_loss=None, ce_loss=None, batch_size: int = 1, num_workers: int = 0, pin_memory: bool = True,
                 smooth: float = 1e-8, multiple_loss: str = "add", **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)
        if multiple_loss != "add":  # "sep" concatenates two 0-d loss tensors in the reference (:90) and cannot run
            raise NotImplementedError("multiple_loss='sep' is broken in the reference (badge_selector.py:90); use 'add'")
        self.dice_smooth, self.dice_do_bg = _dice_params(dice_loss)
        _check_ce(ce_loss)
        self.multiple_loss = multiple_loss

    def cal_scores(self, active_dataset, model, device):
        """(case names, gradient embeddings [n, K*32] on device) of the whole pool, like badge_selector.py:56-104."""
        pool = active_dataset.get_pool_dataset()
        rank, ws = D.world()
        lo, hi = D.shard_range(len(pool), rank, ws)
        src = self._source(pool, device)
        eng = self._engine(model, device, src.image_hw())
        out = torch.empty((hi - lo, 128), dtype=torch.float32, device=eng.device)
        for off, imgs, norm in src.chunks(list(range(lo, hi)), eng.max_samples, eng):
            out[off:off + imgs.shape[0]] = eng.badge_embed(imgs, norm, self.dice_smooth, self.dice_do_bg, 128)
        return list(pool.image_idx), D.gather_rows(out)[:, :eng.K * 32], eng

    def select_next_batch(self, active_dataset, select_num: int, model, device):
        import numpy as np
        labeled_size, pool_size = active_dataset.get_size()
        if labeled_size == 0:
            return _cold_start(active_dataset, select_num)  # badge_selector.py:114-121
        if select_num <= 0:
            return []
        if pool_size < select_num:
            raise ValueError(f"n_samples={pool_size} should be >= n_clusters={select_num}.")
        names, emb, eng = self.cal_scores(active_dataset, model, device)
        X = torch.zeros((pool_size, 128), dtype=torch.float32, device=eng.device)
        X[:, :emb.shape[1]] = emb  # zero columns do not change any distance
        rs = np.random.mtrand._rand
        trials = 2 + int(np.log(select_num))
        w = np.ones(pool_size, dtype=np.float32)  # sample_weight=None
        first = int(rs.choice(pool_size, p=w / w.sum()))
        uniforms = np.stack([rs.uniform(size=trials) for _ in range(select_num - 1)]) if select_num > 1 \
            else np.zeros((0, trials))
        first, uniforms = D.broadcast_draws(first, uniforms, eng.device)
        _require_finite(X, "gradient embeddings")
        picks = eng.kmeans_pp(X, torch.from_numpy(w), first, torch.from_numpy(uniforms), select_num, trials).tolist()
        return [names[i] for i in picks]  # pick order, like list(case_name_list[selected_indices]) (:126)


class B200HybridSelector(_B200Uncertainty):
    """Uncertainty pre-filter (top `prefilter_factor * select_num` by score)
    followed by k-center greedy (l2) on the survivors' embeddings.  Not in the
    reference; composed from its cal_scores and kcenter_greedy."""

    def __init__(self, batch_size: int, num_workers: int, pin_memory: bool = True, smooth: float = 1e-8, *,
                 mode: str = "entropy", prefilter_factor: int = 10, **kw):
        super().__init__(batch_size, num_workers, pin_memory, smooth, **kw)
        self.mode = mode
        self.prefilter_factor = int(prefilter_factor)

    def select_next_batch(self, active_dataset, select_num: int, model, device):
        labeled_size, pool_size = active_dataset.get_size()
        if labeled_size == 0:
            return _cold_start(active_dataset, select_num)
        labeled, pool = active_dataset.get_train_dataset(), active_dataset.get_pool_dataset()
        rank, ws = D.world()
        lo, hi = D.shard_range(pool_size, rank, ws)
        psrc = self._source(pool, device)
        lsrc = self._source(labeled, device)
        eng = self._engine(model, device, psrc.image_hw())
        self._stages.start("score")
        scores = self._score_positions(eng, psrc, list(range(lo, hi)))
        self._stages.start("rank")
        m = min(pool_size, self.prefilter_factor * select_num)
        vals, pos = eng.topk(scores, m, base_pos=lo)
        if ws > 1:
            vals, pos = D.global_topk(vals, pos, m)
        cand = sorted(pos.tolist())  # pool order -> positional tie-break inside k-center
        mine = [p for p in cand if lo <= p < hi]
        self._stages.start("embed")
        emb_mine = _embed_positions(eng, psrc, mine)
        emb_l = _embed_positions(eng, lsrc, list(range(labeled_size)))
        self._stages.start("gather")
        emb_c = D.gather_rows(emb_mine)
        feats = torch.cat([emb_l, emb_c], 0)
        _require_finite(feats, "embeddings")
        self._stages.start("kcenter")
        picks = _check_picks(eng.kcenter(feats, labeled_size, select_num, "l2").tolist(), "k-center")
        return [pool.image_idx[cand[i - labeled_size]] for i in picks]


SELECTORS = {
    "b200-entropy": B200EntropySelector,
    "b200-margin": B200MarginSelector,
    "b200-confidence": B200ConfidenceSelector,
    "b200-dice": B200DiceUncertaintySelector,
    "b200-mc-entropy": B200MCDropoutEntropySelector,
    "b200-coreset-l2": lambda **kw: B200CoresetSelector(metric="l2", **kw),
    "b200-coreset-cosine": lambda **kw: B200CoresetSelector(metric="cosine", **kw),
    "b200-hybrid": B200HybridSelector,
    "b200-kmean-l2": lambda **kw: B200KMeanSelector(metric="l2", **kw),
    "b200-kmean-cosine": lambda **kw: B200KMeanSelector(metric="cosine", **kw),
    "b200-badge": B200BADGESelector,
}
