"""Config-scale goldens from the UNMODIFIED reference (container only; needs /root/reference).

  python -m oracle.make_golden_configs [stage ...]      stages: c1 c2 fwd20k emb100k select  (default: all)

BASELINE.json configs at their stated sizes, frozen into tests/golden/configs.npz:
  c1      config 1: reference Entropy/Margin/Confidence selectors, 512-image BUSI-shape (K=2) pool, 32 labeled,
          budget 32 (entropy_selector.py:24-82)
  c2      config 2: composed MC-dropout oracle (T=8, p=0.1) on the 1000-image prefix of the FUGC-shape (K=3) pool:
          scores + top-100 ("parity unpinned by the reference": MC-dropout does not exist there)
  fwd20k  reference UNet over images 0..20499 of the FUGC-shape set (500 labeled + 20k pool prefix): entropy scores
          (reference formula) and get_enc_feature embeddings from ONE pass - the encoder / decoder halves of
          UNet.forward (unet.py:293) are called separately and checked bit-identical to model(x) /
          model.get_enc_feature(x) on the first batch
  emb100k reference encoder over images 20500..100499 (the rest of the 100k pool)
  select  config 3 at N=20k: the reference's own CoresetSelector.select_next_batch run VERBATIM (its DataLoader,
          sklearn pairwise_distances, /sum, kcenter_greedy, name mapping) over a model whose get_enc_feature returns
          the rows computed above; the streaming oracle must reproduce its set.  Config 3 at 100k: streaming oracle
          on the reference embeddings (the verbatim matrix would need 3 x 40 GB).  Config 4 (hybrid) at 20k and -
          scores permitting - 100k.  Per-step top-2 gaps, names and a sha256 of the embeddings are frozen, not the
          205 MB matrix.

Intermediate arrays live under $B200AL_GOLDEN_TMP (default /tmp/b200al_golden) so stages can be re-run.
Dataset layout of the FUGC-shape set: image i = oracle.synth.synth_image_u8(i); names "%06d"; labeled = the first
500 names, pool position p <-> image 500 + p.
"""
from __future__ import annotations

import hashlib
import os
import sys
import time

import numpy as np
import torch

from . import select_ref as S
from .synth import make_active, synth_pool_u8, case_name
from .unet_ref import golden_state_dict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
TMP = os.environ.get("B200AL_GOLDEN_TMP", "/tmp/b200al_golden")
THREADS = int(os.environ.get("B200AL_GOLDEN_THREADS", "6"))

N_LABELED3 = 500        # configs 3 / 4: labeled rows
N_POOL_SMALL = 20_000   # verbatim reference still fits (1.7 GB per matrix)
N_POOL_FULL = 100_000
BUDGET3 = 500
PREFILTER = 10          # hybrid: top 10 x budget by entropy
MC_T, MC_P, MC_SEED, MC_N, MC_BUDGET = 8, 0.1, 2025, 1000, 100


def log(*a):
    print(time.strftime("%H:%M:%S"), *a, flush=True)


def versions():
    import sklearn
    return dict(torch=torch.__version__, sklearn=sklearn.__version__, numpy=np.__version__, threads=THREADS)


def tmp(name):
    os.makedirs(TMP, exist_ok=True)
    return os.path.join(TMP, name)


def model_for(K):
    from .ref_import import build_reference_unet
    sd = golden_state_dict(K, np.load(os.path.join(OUT, "model.npz")))
    model = build_reference_unet(K)
    model.load_state_dict(sd)
    model.eval()
    return sd, model


# ---------------------------------------------------------------------------------------------------------------
def stage_c1():
    """config 1, verbatim reference selectors on CPU."""
    from .ref_import import load_reference
    ref = load_reference()
    _sd, model = model_for(2)
    images = synth_pool_u8(32 + 512)
    cpu = torch.device("cpu")
    out = {}
    for mode, cls in (("entropy", ref.EntropySelector), ("margin", ref.MarginSelector),
                      ("confidence", ref.ConfidenceSelector)):
        t0 = time.time()
        ad = make_active(images, 32, reference=ref)
        sel = cls(batch_size=8, num_workers=0, pin_memory=False)
        score_list, names = sel.cal_scores(ad, model, cpu)
        scores = torch.stack(score_list).numpy()
        # select_next_batch re-runs cal_scores; rank the frozen scores the reference's way instead (:73-82) and check
        # that the stable tie-break picks the same names
        _, idx = torch.sort(torch.stack(score_list), descending=True)
        picked = [names[i] for i in idx[:32]]
        stable = [names[i] for i in S.rank_desc_stable(scores, 32)]
        assert picked == stable, "reference sort order differs from the stable ranking (tie at the cut?)"
        out[f"c1/scores/{mode}"] = scores
        out[f"c1/selected/{mode}"] = np.array(picked)
        out[f"c1/gap/{mode}"] = np.float64(S.boundary_gap(scores, 32))
        log(f"c1 {mode}: {time.time() - t0:.0f}s gap {out[f'c1/gap/{mode}']:.3e}")
    ad = make_active(images, 32, reference=ref)
    sel = ref.EntropySelector(batch_size=8, num_workers=0, pin_memory=False)
    full = sel.select_next_batch(ad, 32, model, cpu)   # the whole public call, once
    assert list(full) == out["c1/selected/entropy"].tolist()
    np.savez(tmp("c1.npz"), **out)


# ---------------------------------------------------------------------------------------------------------------
def stage_c2():
    """config 2: composed MC-dropout oracle, T=8, 1000-image prefix of the FUGC-shape pool (images 500..1499)."""
    sd, _ = model_for(3)
    path = tmp("c2_scores.npy")
    done = np.load(path) if os.path.exists(path) else np.zeros(0, np.float32)
    step = 40
    while len(done) < MC_N:
        i0 = len(done)
        imgs = synth_pool_u8(step, start=N_LABELED3 + i0)
        pos = np.arange(i0, i0 + step)   # masks are keyed by POOL position
        t0 = time.time()
        s = S.pool_scores_mc(sd, imgs, pos, MC_T, MC_P, MC_SEED, "entropy", batch=4)
        done = np.concatenate([done, s.astype(np.float32)])
        np.save(path, done)
        log(f"c2 {len(done)}/{MC_N} ({(time.time() - t0) / step:.2f} s/img)")
    out = {"c2/scores": done, "c2/selected_pos": S.rank_desc_stable(done, MC_BUDGET).astype(np.int64),
           "c2/gap": np.float64(S.boundary_gap(done, MC_BUDGET))}
    np.savez(tmp("c2.npz"), **out)
    log("c2 gap", out["c2/gap"])


def _reference_split_pass(model, x):
    """(entropy scores, embeddings, logits) of a batch from ONE encoder pass: UNet.forward is
    decoder(encoder(x, return_skips=True)) (unet.py:293) and get_enc_feature pools the last skip (:89-93)."""
    import torch.nn.functional as N
    skips = model.encoder(x, return_skips=True)
    emb = N.adaptive_avg_pool2d(skips[-1], (1, 1)).view(x.shape[0], -1)
    pred = model.decoder(list(skips))
    prob = pred.softmax(1)
    entropy = torch.mean(-prob * torch.log2(prob + 1e-8), dim=1)   # entropy_selector.py:45-49
    return entropy.mean(dim=[-2, -1]), emb, pred


def stage_fwd20k():
    """images 0..20499: entropy scores + embeddings by the reference module."""
    _sd, model = model_for(3)
    n = N_LABELED3 + N_POOL_SMALL
    ps, pe = tmp("fwd20k_scores.npy"), tmp("fwd20k_emb.npy")
    scores = np.load(ps) if os.path.exists(ps) else np.zeros(0, np.float32)
    emb = np.load(pe) if os.path.exists(pe) else np.zeros((0, 512), np.float32)
    assert len(scores) == len(emb)
    bs, chunk = 8, 200
    with torch.no_grad():
        if len(scores) == 0:   # verbatim check on the first batch
            x = synth_pool_u8(bs).float().div(255).unsqueeze(1)
            s, e, pred = _reference_split_pass(model, x)
            assert torch.equal(pred, model(x)) and torch.equal(e, model.get_enc_feature(x))
            log("fwd20k: split pass == model(x) / get_enc_feature(x), bit-identical")
        while len(scores) < n:
            i0 = len(scores)
            m = min(chunk, n - i0)
            imgs = synth_pool_u8(m, start=i0)
            t0 = time.time()
            ss, ee = [], []
            for j in range(0, m, bs):
                x = imgs[j:j + bs].float().div(255).unsqueeze(1)
                s, e, _ = _reference_split_pass(model, x)
                ss.append(s.numpy()); ee.append(e.numpy())
            scores = np.concatenate([scores] + ss).astype(np.float32)
            emb = np.concatenate([emb] + ee).astype(np.float32)
            np.save(ps, scores); np.save(pe, emb)
            log(f"fwd20k {len(scores)}/{n} ({(time.time() - t0) / m * 1e3:.0f} ms/img)")


def stage_emb100k():
    """images 20500..100499: reference get_enc_feature."""
    _sd, model = model_for(3)
    n0 = N_LABELED3 + N_POOL_SMALL
    n = N_LABELED3 + N_POOL_FULL
    pe = tmp("emb100k_rest.npy")
    emb = np.load(pe) if os.path.exists(pe) else np.zeros((0, 512), np.float32)
    bs, chunk = 8, 800
    with torch.no_grad():
        while n0 + len(emb) < n:
            i0 = n0 + len(emb)
            m = min(chunk, n - i0)
            imgs = synth_pool_u8(m, start=i0)
            t0 = time.time()
            ee = [model.get_enc_feature(imgs[j:j + bs].float().div(255).unsqueeze(1)).numpy() for j in range(0, m, bs)]
            emb = np.concatenate([emb] + ee).astype(np.float32)
            np.save(pe, emb)
            log(f"emb100k {n0 + len(emb)}/{n} ({(time.time() - t0) / m * 1e3:.0f} ms/img)")


# ---------------------------------------------------------------------------------------------------------------
class _CachedFeatureModel:
    """Stands in for the UNet inside the VERBATIM CoresetSelector: get_enc_feature returns, batch by batch in
    DataLoader order (labeled rows first, coreset_selector.py:85-97), the embeddings the reference module computed
    for those very images in stage fwd20k / emb100k."""

    def __init__(self, emb):
        self.emb, self.off = torch.from_numpy(np.ascontiguousarray(emb)), 0

    def eval(self):
        return self

    def get_enc_feature(self, x):
        out = self.emb[self.off:self.off + x.shape[0]]
        self.off += x.shape[0]
        return out


def _stream_l2(feats, n_init, budget):
    """oracle.select_ref.kcenter_greedy_stream(l2) with the fp64 copy and row norms hoisted out of the loop
    (identical arithmetic per column: fp64 xx + cc - 2 x.c -> fp32 -> max 0 -> self 0 -> fp32 sqrt)."""
    x64 = feats.astype(np.float64)
    xx = (x64 * x64).sum(1)

    def row_fn(_f, c, j):
        c64 = c.astype(np.float64)
        d2 = np.maximum((xx + (c64 * c64).sum() - 2.0 * (x64 @ c64)).astype(np.float32), np.float32(0))
        d2[j] = 0
        return np.sqrt(d2)

    return S.kcenter_greedy_stream(feats, n_init, budget, "l2", row_fn=row_fn, return_gaps=True)


def _sample_rows(n):
    return np.unique(np.concatenate([np.arange(64), np.arange(0, n, 997)]))


def stage_select():
    from .ref_import import load_reference
    from .synth import SynthBase
    ref = load_reference()
    out = {}
    scores_all = np.load(tmp("fwd20k_scores.npy"))
    emb20 = np.load(tmp("fwd20k_emb.npy"))
    L = N_LABELED3
    assert len(emb20) == L + N_POOL_SMALL
    names20 = [case_name(i) for i in range(L + N_POOL_SMALL)]

    # --- config 3 at 20k: verbatim CoresetSelector.select_next_batch over cached reference embeddings
    class _RefBase(SynthBase, ref.BaseDataset):
        pass
    base = _RefBase(torch.zeros((len(names20), 1, 1), dtype=torch.uint8), names20)   # pixels are not read again
    ad = ref.ActiveDataset(ref.ExtendableDataset(base, names20[:L]), ref.ExtendableDataset(base, names20[L:]))
    t0 = time.time()
    sel = ref.CoresetSelector(batch_size=64, num_workers=0, pin_memory=False, metric="l2")
    verbatim = sel.select_next_batch(ad, BUDGET3, _CachedFeatureModel(emb20), torch.device("cpu"))
    log(f"verbatim CoresetSelector at N={len(names20)}: {time.time() - t0:.0f}s")
    picks, gaps = _stream_l2(emb20, L, BUDGET3)
    assert sorted(map(str, verbatim)) == sorted(names20[q] for q in picks), "streaming oracle != verbatim reference"
    log("coreset 20k: streaming oracle == verbatim reference set; min top-2 gap", min(gaps))
    out["c3_20k/picks"] = np.array(picks, np.int64)     # row indices (labeled rows first), selection order
    out["c3_20k/gaps"] = np.array(gaps, np.float32)
    out["c3_20k/selected"] = np.array(sorted(names20[q] for q in picks))

    # --- config 4 at 20k: hybrid = entropy top-M -> k-center
    pool_scores = scores_all[L:]
    M = PREFILTER * BUDGET3
    hyb = S.hybrid_select(pool_scores, emb20[:L], emb20[L:], BUDGET3, M)
    out["c4_20k/selected_pos"] = np.array(hyb, np.int64)
    out["c4_20k/prefilter_gap"] = np.float64(S.boundary_gap(pool_scores, M))
    keep = np.sort(S.rank_desc_stable(pool_scores, M))
    _p, hgaps = _stream_l2(np.concatenate([emb20[:L], emb20[L:][keep]], 0), L, BUDGET3)
    out["c4_20k/gaps"] = np.array(hgaps, np.float32)
    out["c34/entropy_scores_20k"] = pool_scores.astype(np.float32)
    log("hybrid 20k: prefilter gap", out["c4_20k/prefilter_gap"], "min k-center gap", min(hgaps))

    rows = _sample_rows(len(emb20))
    out["c34/emb_sample_rows"] = rows
    out["c34/emb_sample"] = emb20[rows]
    out["c34/emb20k_sha256"] = np.array(hashlib.sha256(emb20.tobytes()).hexdigest())

    # --- config 3 at 100k: streaming oracle on reference embeddings
    rest = tmp("emb100k_rest.npy")
    if os.path.exists(rest) and len(np.load(rest, mmap_mode="r")) == N_POOL_FULL - N_POOL_SMALL:
        emb = np.concatenate([emb20, np.load(rest)], 0)
        t0 = time.time()
        picks, gaps = _stream_l2(emb, L, BUDGET3)
        log(f"coreset 100k streaming oracle: {time.time() - t0:.0f}s, min top-2 gap {min(gaps):.3e}")
        out["c3_100k/picks"] = np.array(picks, np.int64)
        out["c3_100k/gaps"] = np.array(gaps, np.float32)
        out["c3_100k/selected"] = np.array(sorted(case_name(q) for q in picks))
        rows = _sample_rows(len(emb))
        out["c3_100k/emb_sample_rows"] = rows
        out["c3_100k/emb_sample"] = emb[rows]
        out["c3_100k/emb_sha256"] = np.array(hashlib.sha256(emb.tobytes()).hexdigest())
    else:
        log("emb100k not complete: config 3 at 100k skipped")
    np.savez(tmp("select.npz"), **out)


KM_SEED = 4242


def stage_kmean():
    """KMeanSelector at config-3 scale: the reference's own select_next_batch VERBATIM (its
    DataLoaders - pool first, then labeled, kmean_selector.py:146-156 -, row standardisation, sklearn
    pairwise_distances + kmeans_plusplus on numpy's global RNG) over the cached reference embeddings of the 20 500
    images, cosine (its default metric) and l2; frozen: the selected names, the reference's and the oracle's pick
    order, the pick at which they part (if they do) and the oracle's conditioning."""
    from .kmean_ref import kmean_select
    from .ref_import import load_reference
    from .synth import SynthBase
    ref = load_reference()
    out = {}
    emb20 = np.load(tmp("fwd20k_emb.npy"))
    L = N_LABELED3
    names20 = [case_name(i) for i in range(L + N_POOL_SMALL)]

    class _RefBase(SynthBase, ref.BaseDataset):
        pass
    base = _RefBase(torch.zeros((len(names20), 1, 1), dtype=torch.uint8), names20)
    for metric in ("cosine", "l2"):
        ad = ref.ActiveDataset(ref.ExtendableDataset(base, names20[:L]), ref.ExtendableDataset(base, names20[L:]))
        sel = ref.KMeanSelector(batch_size=64, num_workers=0, pin_memory=False, metric=metric)
        np.random.seed(KM_SEED)
        t0 = time.time()
        served = _CachedFeatureModel(np.concatenate([emb20[L:], emb20[:L]], 0))   # pool rows are asked for first
        names = sel.select_next_batch(ad, BUDGET3, served, torch.device("cpu"))
        log(f"verbatim KMeanSelector {metric} at N={len(names20)}: {time.time() - t0:.0f}s, {len(names)} names")
        want = sorted(map(str, names))
        # the same call with the picks in order (the selector returns an unordered set, kmean_selector.py:194)
        from sklearn.cluster import kmeans_plusplus
        from .kmean_ref import sample_weights, standardize_rows
        pool_std, lab_std = standardize_rows(emb20[L:]), standardize_rows(emb20[:L])
        w = sample_weights(pool_std, lab_std, metric, "min", 1.0, False)
        _c, ref_idx = kmeans_plusplus(pool_std, BUDGET3, sample_weight=w, random_state=np.random.RandomState(KM_SEED))
        # (not asserted equal to the verbatim names: the selector's weights come from sklearn's fp32 sgemm cosine /
        # fp64 l2 and differ from the restatement's in the last ulp on some of the 20 000 rows - enough, at this size,
        # to move a uniform draw across a cumulative-weight boundary; see the comment below)
        idx, gap = kmean_select(emb20[L:], emb20[:L], BUDGET3, metric, "min", 1.0, False,
                                np.random.RandomState(KM_SEED), return_gap=True)
        # sklearn carries the winning potential as an fp32 BLAS dot (machine-specific summation order); the oracle and
        # the GPU kernel sum it in fp64 and round.  One ulp of difference moves the next step's uniform draws by 6e-8
        # relative, and with 20 000 cumulative-weight boundaries and 4 000 draws a run has a draw that close to a
        # boundary about once: from there on the two runs sample different candidates.
        div = next((i for i, (a, b) in enumerate(zip(idx, ref_idx)) if int(a) != int(b)), -1)
        out[f"km_20k/{metric}/selected"] = np.array(want)
        out[f"km_20k/{metric}/reference_picks"] = np.array([int(i) for i in ref_idx], np.int64)  # pool positions
        out[f"km_20k/{metric}/oracle_picks"] = np.array([int(i) for i in idx], np.int64)
        out[f"km_20k/{metric}/first_divergence"] = np.int64(div)
        out[f"km_20k/{metric}/gap"] = np.float64(gap)

        def potential(pos):   # the k-means++ objective of a selected set: sum_i w_i min_c |x_i - c|^2
            x64, c64 = pool_std.astype(np.float64), pool_std[np.asarray(pos)].astype(np.float64)
            d2 = np.full(len(x64), np.inf)
            for a in range(0, len(c64), 50):
                cc = c64[a:a + 50]
                d2 = np.minimum(d2, ((x64 * x64).sum(1)[:, None] + (cc * cc).sum(1)[None] - 2 * x64 @ cc.T).min(1))
            return float((np.maximum(d2, 0) * w.astype(np.float64)).sum())
        verb_pos = [int(n) - L for n in want]     # case names are zero-padded global indices
        rs2 = np.random.RandomState(KM_SEED + 1)
        _c, other = kmeans_plusplus(pool_std, BUDGET3, sample_weight=w, random_state=rs2)   # another seed: the spread
        pots = {"verbatim": potential(verb_pos), "oracle": potential(idx), "sklearn_on_oracle_features": potential(ref_idx),
                "other_seed": potential(other)}
        for k_, v_ in pots.items():
            out[f"km_20k/{metric}/potential/{k_}"] = np.float64(v_)
        log(f"kmean {metric}: verbatim-vs-oracle overlap {len(set(verb_pos) & set(map(int, idx)))}/{BUDGET3}; potentials {pots}")
        log(f"kmean {metric}: oracle vs verbatim reference: first divergence at pick {div} (-1 = none), overlap "
            f"{len(set(map(int, idx)) & set(map(int, ref_idx)))}/{BUDGET3}, conditioning gap {gap:.3e}")
    out["km_20k/np_seed"] = np.int64(KM_SEED)
    np.savez(tmp("kmean.npz"), **out)


def merge():
    path = os.path.join(OUT, "configs.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}   # stages frozen by earlier runs stay
    out["versions"] = np.array(repr(versions()))
    for f in ("c1.npz", "c2.npz", "select.npz", "kmean.npz"):
        if os.path.exists(tmp(f)):
            out.update(np.load(tmp(f)))
    np.savez_compressed(path, **out)
    log("wrote", path, os.path.getsize(path), "bytes:", sorted(out))


STAGES = {"c1": stage_c1, "c2": stage_c2, "fwd20k": stage_fwd20k, "emb100k": stage_emb100k, "select": stage_select,
          "kmean": stage_kmean}


def main(argv):
    if not os.path.isdir("/root/reference"):
        sys.exit("needs /root/reference")
    torch.set_num_threads(THREADS)
    for s in (argv or list(STAGES)):
        log("stage", s)
        STAGES[s]()
        merge()


if __name__ == "__main__":
    main(sys.argv[1:])
