"""CPU restatement of the reference's KMeanSelector (oracle side; TEST INFRASTRUCTURE ONLY).

Path restated (reference file:line):
  * `KMeanSelector.get_features` (src/activelearning/kmean_selector.py:48-127): `model.get_enc_feature` embeddings,
    then per-ROW standardisation `(feats - mean(axis=1)) / std(axis=1)` in fp32 numpy (:95-97).
  * `cal_scores` (:129-161): `pairwise_distances(pool_feats, labeled_feats, metric)` (sklearn fp32 semantics, same as
    oracle/select_ref.py's distance functions).
  * `select_next_batch` (:163-196): `sample_weight` = row min (or mean) of that matrix, `** sharp_factor / sum` or
    `softmax(sample_weight * sharp_factor)` (:175-186); `kmeans_plusplus(X=pool_feats, n_clusters=select_num,
    sample_weight=...)` with the GLOBAL numpy RNG (:190-192); `list(set(names))` (:194).  With no labeled data the
    weights are None -> uniform (:187-188).
  * `sklearn.cluster.kmeans_plusplus` is third-party (scikit-learn, unpinned in pyproject.toml:25; 1.9.0 installed
    here and on the GPU box).  Its `_kmeans_plusplus` (sklearn/cluster/_kmeans.py) is restated from the installed
    source: first centre `random_state.choice(n, p=w/w.sum())`; then per centre `n_local_trials = 2 + int(log(k))`
    candidates at `searchsorted(cumsum_f32(w * closest_sq), uniform(size=trials) * current_pot)`, squared euclidean
    distances `f32(max(xx + yy - 2xy, 0))` evaluated in fp64, `minimum` with the running closest distances, the
    candidate with the smallest weighted potential wins (first on ties).
    One deliberate difference: sklearn evaluates the potentials `dist @ w` with an fp32 BLAS dot whose summation
    order is CPU-specific; here (and in the CUDA kernels) they are accumulated in fp64 and rounded to fp32.  The two
    agree to ~1e-7 relative, so a pick can only differ when a uniform draw lands within that distance of a
    cumulative-weight boundary (probability ~ n * 1e-7 per draw); tests pin index-exact agreement with sklearn itself
    on seeded pools.
"""
from __future__ import annotations

import numpy as np

from .select_ref import cosine_rows, l2_rows_direct, normalize_rows


def standardize_rows(feats: np.ndarray) -> np.ndarray:
    """kmean_selector.py:95-97 (np.std: population std, ddof=0), fp32."""
    feats = np.asarray(feats, np.float32)
    mean = np.mean(feats, axis=1, keepdims=True)
    std = np.std(feats, axis=1, keepdims=True)
    return ((feats - mean) / std).astype(np.float32)


def sq_dist_f32(c: np.ndarray, X: np.ndarray) -> np.ndarray:
    """sklearn `_euclidean_distances(..., squared=True)` for fp32 inputs: fp64 evaluation, fp32 result, clamped at 0."""
    c64, X64 = c.astype(np.float64), X.astype(np.float64)
    d = ((X64[None, :, :] - c64[:, None, :]) ** 2).sum(-1)
    return np.maximum(d.astype(np.float32), np.float32(0))


def sample_weights(pool_feats, labeled_feats, metric="cosine", criteria="min", sharp_factor=1.0, softmax=False):
    """kmean_selector.py:143-145,174-186 -> normalised fp32 weights, or None without labeled data."""
    if labeled_feats is None or len(labeled_feats) == 0:
        return None
    if metric == "l2":
        dist = np.stack([l2_rows_direct(pool_feats, c) for c in labeled_feats], axis=1)
    else:
        pn, ln = normalize_rows(pool_feats), normalize_rows(labeled_feats)
        dist = np.stack([cosine_rows(pn, c) for c in ln], axis=1)
    w = dist.min(axis=1) if criteria == "min" else dist.astype(np.float64).mean(axis=1).astype(np.float32)
    return normalise_weights(w, sharp_factor, softmax)


def normalise_weights(w: np.ndarray, sharp_factor=1.0, softmax=False) -> np.ndarray:
    if softmax:
        import torch
        return (torch.from_numpy(np.asarray(w)) * sharp_factor).softmax(0).numpy()
    w = w ** sharp_factor
    return w / w.sum()


def kmeans_plusplus_indices(X: np.ndarray, n_clusters: int, sample_weight: np.ndarray | None, random_state,
                            return_gap: bool = False):
    """Indices chosen by sklearn 1.9.0 `_kmeans_plusplus` (potentials in fp64 -> fp32, see module docstring).
    return_gap: also the conditioning of the run = the smallest relative distance of any uniform draw to a
    cumulative-weight boundary and of any winning potential to the runner-up; perturbations of the inputs well below
    it (e.g. bf16x3 embeddings instead of fp32 ones) cannot change the picks."""
    X = np.asarray(X, np.float32)
    n = X.shape[0]
    w = np.ones(n, np.float32) if sample_weight is None else np.asarray(sample_weight, np.float32)
    trials = 2 + int(np.log(n_clusters))
    idx = np.full(n_clusters, -1, dtype=np.int64)
    idx[0] = random_state.choice(n, p=w / w.sum())
    closest = sq_dist_f32(X[idx[0]][None], X)[0]
    pot = np.float32((closest.astype(np.float64) * w.astype(np.float64)).sum())
    gap = np.inf
    for c in range(1, n_clusters):
        rand_vals = random_state.uniform(size=trials) * pot
        cum = np.cumsum(w * closest)
        cand = np.searchsorted(cum, rand_vals)
        np.clip(cand, None, n - 1, out=cand)
        d = np.minimum(closest[None, :], sq_dist_f32(X[cand], X))
        pots = (d.astype(np.float64) @ w.astype(np.float64)).astype(np.float32)
        best = int(np.argmin(pots))
        if return_gap and pot > 0:
            lower = np.where(cand > 0, cum[np.maximum(cand - 1, 0)], 0.0)
            gap = min(gap, float(np.min(np.minimum(rand_vals - lower, cum[cand] - rand_vals)) / pot))
            others = pots[cand != cand[best]]
            if others.size:
                gap = min(gap, float((others.min() - pots[best]) / pot))
        pot, closest, idx[c] = pots[best], d[best], cand[best]
    return (idx, gap) if return_gap else idx


def kmean_select(pool_emb, labeled_emb, select_num, metric="cosine", criteria="min", sharp_factor=1.0, softmax=False,
                 random_state=None, return_gap: bool = False):
    """Pool positions picked by KMeanSelector.select_next_batch, in pick order (the reference returns them as an
    unordered `list(set(...))`, :194).  Embeddings are the raw `get_enc_feature` rows."""
    rs = np.random.mtrand._rand if random_state is None else random_state
    pool = standardize_rows(pool_emb)
    lab = standardize_rows(labeled_emb) if labeled_emb is not None and len(labeled_emb) else None
    w = sample_weights(pool, lab, metric, criteria, sharp_factor, softmax)
    return kmeans_plusplus_indices(pool, select_num, w, rs, return_gap)


def kmean_features(emb, loaded, weight: float = 1.0, numpy_rule: str = "legacy") -> np.ndarray:
    """KMeanSelector.get_features after the loops (kmean_selector.py:92-125): [standardised embeddings |
    standardised loaded rows * np.sqrt(D1 / D2 * loaded_feature_weight)]; either part may be None
    (`loaded_feature_only`, or no loaded features).  numpy_rule: see oracle.select_ref.cat_scale."""
    from .select_ref import cat_scale
    parts = []
    if emb is not None:
        parts.append(standardize_rows(emb))
    if loaded is not None:
        lf = standardize_rows(loaded)
        parts.append(lf if emb is None else lf * cat_scale(np.shape(emb)[1], lf.shape[1], weight, numpy_rule))
    return np.concatenate(parts, axis=1)


def kmean_mix_select(pool_emb, labeled_emb, pool_loaded, labeled_loaded, weight, select_num, metric="cosine",
                     criteria="min", sharp_factor=1.0, softmax=False, random_state=None, return_gap=False):
    """KMeanSelector.select_next_batch with model AND loaded features under the reference's numpy < 2 rule (float32
    throughout): pool positions in pick order."""
    rs = np.random.mtrand._rand if random_state is None else random_state
    pool = kmean_features(pool_emb, pool_loaded, weight)
    lab = kmean_features(labeled_emb, labeled_loaded, weight) if labeled_emb is not None and len(labeled_emb) else None
    w = sample_weights(pool, lab, metric, criteria, sharp_factor, softmax)
    return kmeans_plusplus_indices(pool, select_num, w, rs, return_gap)
