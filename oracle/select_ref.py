"""Restatement of the reference's score / rank / k-center logic (oracle; test-only).

Each function cites the reference lines it follows.  The distance formulas are
those of scikit-learn 1.9.0 (`pyproject.toml:33` pins only `scikit-learn>=1.3`;
1.9.0 is what the reference resolves to in this image), which the reference
calls at src/activelearning/coreset_selector.py:132,152.
"""
from __future__ import annotations

import numpy as np
import torch

SMOOTH = 1e-8  # EntropySelector default (entropy_selector.py:16)


# ---------------------------------------------------------------------------
# per-image uncertainty scores from logits [B,K,H,W]
# ---------------------------------------------------------------------------
def entropy_scores(logits: torch.Tensor, smooth: float = SMOOTH) -> torch.Tensor:
    """entropy_selector.py:45-49: softmax -> mean_k(-p*log2(p+smooth)) -> mean_hw."""
    prob = logits.softmax(1)
    ent = torch.mean(-prob * torch.log2(prob + smooth), dim=1)
    return ent.mean(dim=[-2, -1])


def margin_scores(logits: torch.Tensor) -> torch.Tensor:
    """margin_selector.py:45-48: -(p_top1 - p_top2), mean over pixels."""
    prob = logits.softmax(1)
    top2 = prob.topk(2, dim=1)[0]
    return (-1 * (top2[:, 0] - top2[:, 1])).mean(dim=[-2, -1])


def confidence_scores(logits: torch.Tensor) -> torch.Tensor:
    """confidence_selector.py:45-47: -max_k p, mean over pixels."""
    prob = logits.softmax(1)
    return (-1 * prob.max(1)[0]).mean(dim=[-2, -1])


def scores_from_prob(prob: torch.Tensor, mode: str, smooth: float = SMOOTH) -> torch.Tensor:
    """Same three formulas applied to a probability map (used by the composed
    MC-dropout oracle on the T-averaged probabilities)."""
    if mode == "entropy":
        return torch.mean(-prob * torch.log2(prob + smooth), dim=1).mean(dim=[-2, -1])
    if mode == "margin":
        top2 = prob.topk(2, dim=1)[0]
        return (-1 * (top2[:, 0] - top2[:, 1])).mean(dim=[-2, -1])
    if mode == "confidence":
        return (-1 * prob.max(1)[0]).mean(dim=[-2, -1])
    if mode == "dice":
        return dice_scores_from_prob(prob)
    raise ValueError(mode)


DICE_SMOOTH = 1e-5  # DiceLoss(smooth) as al_train builds it (al_trainer.py:785-791)


def dice_scores_from_prob(prob: torch.Tensor, smooth: float = DICE_SMOOTH) -> torch.Tensor:
    """Dice-uncertainty (named by BASELINE.json's north star, absent from the reference's selectors):
    per image, DiceLoss of the probability map against its own hard prediction, following DiceLoss.forward
    (src/losses/dice_loss.py:32-72) with do_bg=True, batch=False, squared=False: one-hot of argmax (:24-30),
    intersect / sum_inputs / sum_targets over the pixel axes (:62-64), `1 - (2I + s) / (P + Y + s)` per class (:71),
    mean over classes (:75) - evaluated per image (B = 1) so that each image gets its own score."""
    y = torch.zeros_like(prob).scatter_(1, prob.argmax(1, keepdim=True), 1.0)
    axes = tuple(range(2, prob.ndim))
    intersect = (prob * y).sum(axes)
    dice = 1 - (2 * intersect + smooth) / (prob.sum(axes) + y.sum(axes) + smooth)
    return dice.mean(1)


def dice_scores(logits: torch.Tensor, smooth: float = DICE_SMOOTH) -> torch.Tensor:
    return dice_scores_from_prob(logits.softmax(1), smooth)


SCORE_FNS = {"entropy": entropy_scores, "margin": margin_scores, "confidence": confidence_scores,
             "dice": dice_scores}


def rank_desc_stable(scores, k: int) -> np.ndarray:
    """entropy_selector.py:76-82 with the tie-break made explicit: descending
    score, lower pool position first (the reference's torch.sort is unstable,
    so its tie order is implementation-defined)."""
    s = torch.as_tensor(np.asarray(scores, dtype=np.float32))
    _, idx = torch.sort(s, descending=True, stable=True)
    return idx[:k].numpy()


def boundary_gap(scores, k: int) -> float:
    """score_k - score_{k+1} of the descending ranking (0 if k >= n)."""
    s = np.sort(np.asarray(scores, dtype=np.float64))[::-1]
    if k >= len(s) or k <= 0:
        return float("inf")
    return float(s[k - 1] - s[k])


# ---------------------------------------------------------------------------
# scikit-learn 1.9.0 distance formulas (metrics/pairwise.py)
# ---------------------------------------------------------------------------
def l2_rows(feats: np.ndarray, centre: np.ndarray, self_index: int | None = None) -> np.ndarray:
    """Distance of every row of `feats` (fp32) to `centre` (fp32), the way
    `pairwise_distances(feats, metric="l2")` produces one column:
    euclidean_distances -> _euclidean_distances_upcast (pairwise.py:567-638):
    fp64 `xx + yy - 2 x.y`, cast to fp32, max(.,0), self distance forced to 0,
    then sqrt IN fp32 (pairwise.py:403-426)."""
    x = feats.astype(np.float64)
    c = centre.astype(np.float64)
    d2 = (x * x).sum(1) + (c * c).sum() - 2.0 * (x @ c)
    d2 = np.maximum(d2.astype(np.float32), np.float32(0))
    if self_index is not None:
        d2[self_index] = 0
    return np.sqrt(d2)


def l2_rows_direct(feats: np.ndarray, centre: np.ndarray, self_index: int | None = None) -> np.ndarray:
    """What the CUDA kernel evaluates: sum_k (x_k - c_k)^2 accumulated in fp64,
    rounded to fp32, sqrt in fp32.  Equal to `l2_rows` except where the fp64
    value lies within ~1e-14 relative of an fp32 rounding boundary."""
    diff = feats.astype(np.float64) - centre.astype(np.float64)[None, :]
    d2 = (diff * diff).sum(1).astype(np.float32)
    if self_index is not None:
        d2[self_index] = 0
    return np.sqrt(d2)


def normalize_rows(feats: np.ndarray) -> np.ndarray:
    """sklearn.preprocessing.normalize(X) as cosine_similarity applies it
    (fp32 in, fp32 out; zero rows stay zero)."""
    norms = np.sqrt((feats.astype(np.float32) ** 2).sum(1, dtype=np.float32))
    norms[norms == 0] = 1
    return feats / norms[:, None]


def cosine_rows(feats_n: np.ndarray, centre_n: np.ndarray, self_index: int | None = None) -> np.ndarray:
    """cosine_distances (pairwise.py:1171-1182): clip(1 - x^.y^, 0, 2), self 0.
    The reference gets the dot products from an fp32 BLAS GEMM whose summation
    order is unspecified, so this is reproducible only to fp32 rounding."""
    s = (feats_n.astype(np.float32) @ centre_n.astype(np.float32)).astype(np.float32)
    d = np.clip(np.float32(1) - s, np.float32(0), np.float32(2))
    if self_index is not None:
        d[self_index] = 0
    return d


# ---------------------------------------------------------------------------
# k-center greedy
# ---------------------------------------------------------------------------
def kcenter_greedy_matrix(dist_mat: np.ndarray, budget: int, init_idx, criteria: str = "min") -> list[int]:
    """coreset_selector.py:19-51 restated on a full distance matrix, returning
    the picks IN SELECTION ORDER (the reference returns
    `list(set(selected) - set(init_idx))`, order lost - line 49)."""
    n = dist_mat.shape[0]
    labeled = np.zeros(n, dtype=bool)
    labeled[np.asarray(init_idx)] = True
    picks = []
    for _ in range(budget):
        sub = dist_mat[~labeled][:, labeled]             # :31
        nearest = sub.min(axis=1) if criteria == "min" else sub.mean(axis=1)   # :34 / :36
        q = np.arange(n)[~labeled][nearest.argmax()]     # :43-44 (first max)
        labeled[q] = True                                # :45
        picks.append(int(q))
    return picks


def kcenter_greedy_mean_stream(feats: np.ndarray, n_init: int, budget: int, metric: str = "l2",
                               return_gaps: bool = False):
    """criteria "mean" (coreset_selector.py:35-36) without the N x N matrix: the
    score of an unlabeled row is its mean distance to all current centres, kept
    as a running fp64 sum of the fp32 distances (the reference averages an fp32
    matrix slice with numpy's pairwise fp32 summation, so the two agree to fp32
    rounding of the mean; `return_gaps` exposes the top-2 margin per step)."""
    feats = np.ascontiguousarray(feats, dtype=np.float32)
    n = feats.shape[0]
    fn, f = (l2_rows, feats) if metric == "l2" else (cosine_rows, normalize_rows(feats))
    msum = np.zeros(n, dtype=np.float64)
    for j in range(n_init):
        msum += fn(f, f[j], j).astype(np.float64)
    taken = np.zeros(n, dtype=bool)
    taken[:n_init] = True
    picks, gaps = [], []
    for step in range(budget):
        mean = (msum / float(n_init + step)).astype(np.float32)
        cand = np.where(taken, -np.inf, mean)
        q = int(cand.argmax())
        if return_gaps:
            top = cand[q]
            cand[q] = -np.inf
            gaps.append(float(top - cand.max()) if n - taken.sum() > 1 else float("inf"))
        taken[q] = True
        picks.append(q)
        msum += fn(f, f[q], q).astype(np.float64)
    return (picks, gaps) if return_gaps else picks


def kcenter_greedy_stream(feats: np.ndarray, n_init: int, budget: int, metric: str = "l2",
                          row_fn=None, return_gaps: bool = False):
    """Streaming restatement: rows 0..n_init-1 are the labeled set (init
    centres, coreset_selector.py:85,226); `mind[i] = min_j d(i, centre_j)` is
    kept as a running minimum, which equals line 34's min over the growing
    labeled axis.  argmax takes the FIRST maximum in ascending index order
    (numpy argmax, line 43).  The reference's global `feat_dist /= sum`
    (line 153) is a positive scale and is omitted (argmax-invariant up to
    1-ulp quotient ties)."""
    feats = np.ascontiguousarray(feats, dtype=np.float32)
    n = feats.shape[0]
    if metric == "l2":
        fn = row_fn or l2_rows
        f = feats
    elif metric == "cosine":
        fn = row_fn or cosine_rows
        f = normalize_rows(feats)
    else:
        raise ValueError(metric)
    mind = np.full(n, np.inf, dtype=np.float32)
    for j in range(n_init):
        mind = np.minimum(mind, fn(f, f[j], j))
    taken = np.zeros(n, dtype=bool)
    taken[:n_init] = True
    picks, gaps = [], []
    for _ in range(budget):
        cand = np.where(taken, -np.inf, mind)
        q = int(cand.argmax())
        if return_gaps:
            top = cand[q]
            cand[q] = -np.inf
            second = cand.max() if n - taken.sum() > 1 else -np.inf
            gaps.append(float(top - second))
        taken[q] = True
        picks.append(q)
        mind = np.minimum(mind, fn(f, f[q], q))
    return (picks, gaps) if return_gaps else picks


# ---------------------------------------------------------------------------
# whole-pool oracles on top of oracle.unet_ref (same arithmetic as the
# reference selectors' cal_scores loops; batch size does not change results
# in eval mode because BatchNorm uses running statistics)
# ---------------------------------------------------------------------------
def pool_scores(sd, images_u8: torch.Tensor, mode: str = "entropy", batch: int = 8) -> np.ndarray:
    """EntropySelector/MarginSelector/ConfidenceSelector.cal_scores
    (entropy_selector.py:24-54): one fp32 score per pool image, pool order."""
    from .unet_ref import unet_forward
    out = []
    with torch.no_grad():
        for i in range(0, images_u8.shape[0], batch):
            x = images_u8[i:i + batch].to(torch.float32).div(255.0).unsqueeze(1)
            out.append(SCORE_FNS[mode](unet_forward(sd, x)))
    return torch.cat(out).numpy()


def pool_embeddings(sd, images_u8: torch.Tensor, batch: int = 8) -> np.ndarray:
    """CoresetSelector.cal_scores feature loop (coreset_selector.py:99-108)."""
    from .unet_ref import unet_embed
    out = []
    with torch.no_grad():
        for i in range(0, images_u8.shape[0], batch):
            x = images_u8[i:i + batch].to(torch.float32).div(255.0).unsqueeze(1)
            out.append(unet_embed(sd, x))
    return torch.cat(out).numpy()


def pool_scores_mc(sd, images_u8: torch.Tensor, positions, T: int, p: float, seed: int,
                   mode: str = "entropy", batch: int = 4) -> np.ndarray:
    """COMPOSED oracle (not in the reference; parity unpinned by it): reference
    UNet in eval mode with every Dropout2d replaced by an explicit keep mask
    m/(1-p) (oracle.synth.mc_masks), T passes, probabilities averaged over T,
    then the reference score formula on the mean probability map."""
    from .synth import mc_masks
    from .unet_ref import unet_forward
    out = []
    with torch.no_grad():
        for i in range(0, images_u8.shape[0], batch):
            x = images_u8[i:i + batch].to(torch.float32).div(255.0).unsqueeze(1)
            masks = mc_masks(seed, positions[i:i + batch], T, p)   # [b,T,2944]
            prob = 0
            for t in range(T):
                prob = prob + unet_forward(sd, x, masks=masks[:, t]).softmax(1)
            out.append(scores_from_prob(prob / T, mode))
    return torch.cat(out).numpy()


def hybrid_select(scores: np.ndarray, feats_labeled: np.ndarray, feats_pool: np.ndarray,
                  budget: int, prefilter: int) -> list[int]:
    """COMPOSED oracle: keep the top-`prefilter` pool positions by score
    (stable, lower position first), then k-center greedy (l2, min) over
    labeled + kept embeddings; returns pool positions in selection order."""
    keep = rank_desc_stable(scores, prefilter)
    keep_sorted = np.sort(keep)  # candidates stay in pool order -> positional tie-break
    feats = np.concatenate([feats_labeled, feats_pool[keep_sorted]], 0)
    picks = kcenter_greedy_stream(feats, feats_labeled.shape[0], budget, "l2")
    return [int(keep_sorted[q - feats_labeled.shape[0]]) for q in picks]


# ---------------------------------------------------------------------------
# coreset with externally loaded features, coreset_fusion="add" (coreset_selector.py:110-157)
# ---------------------------------------------------------------------------
def pair_distance_sum(feats: np.ndarray, metric: str = "l2") -> float:
    """`pairwise_distances(feats, metric).sum()` (coreset_selector.py:144-145,152-153) accumulated in fp64 row by row
    (the reference sums an fp32 matrix with numpy's pairwise fp32 summation: equal to ~1e-6 relative)."""
    feats = np.ascontiguousarray(feats, dtype=np.float32)
    fn, f = (l2_rows, feats) if metric == "l2" else (cosine_rows, normalize_rows(feats))
    return float(sum(fn(f, f[j], j).astype(np.float64).sum() for j in range(f.shape[0])))


def kcenter_fused_add_stream(feats: np.ndarray, loaded: np.ndarray, weight: float, n_init: int, budget: int,
                             metric: str = "l2", criteria: str = "min", return_gaps: bool = False):
    """k-center greedy on `final = w * D_loaded / sum(D_loaded) + (1 - w) * D_feat / sum(D_feat)`
    (coreset_selector.py:138-157, fp32 element arithmetic in the reference's order) without building the matrices:
    running min (fp32) or running sum (fp64) of the fused distance to the growing centre set."""
    feats = np.ascontiguousarray(feats, dtype=np.float32)
    loaded = np.ascontiguousarray(loaded, dtype=np.float32)
    n = feats.shape[0]
    fn, f, g = (l2_rows, feats, loaded) if metric == "l2" else (cosine_rows, normalize_rows(feats), normalize_rows(loaded))
    s_feat = np.float32(pair_distance_sum(feats, metric))
    s_load = np.float32(pair_distance_sum(loaded, metric))
    w_l, w_f = np.float32(weight), np.float32(1 - weight)

    def fused(j):
        return (w_l * (fn(g, g[j], j) / s_load) + w_f * (fn(f, f[j], j) / s_feat)).astype(np.float32)

    state = np.zeros(n, np.float64) if criteria == "mean" else np.full(n, np.inf, np.float32)
    for j in range(n_init):
        state = state + fused(j) if criteria == "mean" else np.minimum(state, fused(j))
    taken = np.zeros(n, dtype=bool)
    taken[:n_init] = True
    picks, gaps = [], []
    for step in range(budget):
        val = (state / float(n_init + step)).astype(np.float32) if criteria == "mean" else state
        cand = np.where(taken, -np.inf, val)
        q = int(cand.argmax())
        if return_gaps:
            top = cand[q]
            cand[q] = -np.inf
            gaps.append(float((top - cand.max()) / top) if n - taken.sum() > 1 else float("inf"))
        taken[q] = True
        picks.append(q)
        state = state + fused(q) if criteria == "mean" else np.minimum(state, fused(q))
    return (picks, gaps) if return_gaps else picks


# ---------------------------------------------------------------------------
# coreset_fusion="cat" (coreset_selector.py:158-175) / KMeanSelector's model + loaded mix (kmean_selector.py:106-121)
# ---------------------------------------------------------------------------
def cat_scale(d1: int, d2: int, weight: float, numpy_rule: str = "legacy"):
    """`scale_factor = np.sqrt(D1 / D2 * loaded_feature_weight)` - an np.float64 scalar - as the factor the reference's
    `loaded_feats * scale_factor` multiplies with.  The reference pins numpy < 2 (pyproject.toml:21): value-based
    casting keeps `float32_array * float64_scalar` in float32, the scalar is cast to float32 first ("legacy", what the
    product path implements).  Under numpy >= 2 (NEP 50; 2.3.5 is what this image has) the same line promotes the
    array to float64 ("nep50"; used only to pin the restatement against the reference run live here)."""
    s = np.sqrt(d1 / d2 * weight)
    return np.float32(s) if numpy_rule == "legacy" else np.float64(s)


def coreset_cat_rows(feats: np.ndarray, loaded: np.ndarray, weight: float, numpy_rule: str = "legacy") -> np.ndarray:
    """final_feats of the "cat" branch (:160-174): [feats | loaded * scale]; float32 under the legacy rule, float64
    under NEP 50 (np.concatenate promotes)."""
    feats = np.asarray(feats, np.float32)
    loaded = np.asarray(loaded, np.float32)
    return np.concatenate([feats, loaded * cat_scale(feats.shape[1], loaded.shape[1], weight, numpy_rule)], axis=1)


def coreset_cat_select(feats, loaded, weight, n_init, budget, metric="l2", criteria="min", numpy_rule="legacy",
                       return_gaps=False):
    """k-center greedy over the concatenated rows (:175, then :221-229).  Under "nep50" the rows are float64 and
    sklearn's euclidean_distances skips the float32 upcast path: plain fp64 distances, fp64 sqrt."""
    rows = coreset_cat_rows(feats, loaded, weight, numpy_rule)
    if rows.dtype == np.float64:
        def l2_64(f, c, j):
            x = f.astype(np.float64)
            d2 = np.maximum((x * x).sum(1) + (c * c).sum() - 2.0 * (x @ c), 0)
            d2[j] = 0
            return np.sqrt(d2)

        def cos_64(f, c, j):
            d = np.clip(1.0 - f @ c, 0, 2)
            d[j] = 0
            return d
        if metric == "l2":
            f, fn = rows, l2_64
        else:
            nr = np.sqrt((rows * rows).sum(1))
            nr[nr == 0] = 1
            f, fn = rows / nr[:, None], cos_64
        return _kcenter_generic(f, n_init, budget, fn, criteria, return_gaps)
    if criteria == "mean":
        return kcenter_greedy_mean_stream(rows, n_init, budget, metric, return_gaps)
    return kcenter_greedy_stream(rows, n_init, budget, metric, return_gaps=return_gaps)


def _kcenter_generic(f, n_init, budget, fn, criteria, return_gaps):
    n = f.shape[0]
    state = np.zeros(n) if criteria == "mean" else np.full(n, np.inf)
    for j in range(n_init):
        d = fn(f, f[j], j)
        state = state + d if criteria == "mean" else np.minimum(state, d)
    taken = np.zeros(n, dtype=bool)
    taken[:n_init] = True
    picks, gaps = [], []
    for step in range(budget):
        val = state / float(n_init + step) if criteria == "mean" else state
        cand = np.where(taken, -np.inf, val)
        q = int(cand.argmax())
        if return_gaps:
            top = cand[q]
            cand[q] = -np.inf
            gaps.append(float(top - cand.max()) if n - taken.sum() > 1 else float("inf"))
        taken[q] = True
        picks.append(q)
        d = fn(f, f[q], q)
        state = state + d if criteria == "mean" else np.minimum(state, d)
    return (picks, gaps) if return_gaps else picks
