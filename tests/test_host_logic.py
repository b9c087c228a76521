"""Host-side logic and the C-ABI surface, no GPU needed."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import b200al
from b200al import _lib
from b200al import dist as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "b200al.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200al_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_header_symbol():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200al.h but not exported"
        assert s in _lib.PROTOTYPES, f"{s} has no ctypes prototype"
    assert sorted(_lib.PROTOTYPES) == syms


def test_weights_struct_layout():
    assert C.sizeof(_lib.BlockWeights) == 6 * 8
    assert C.sizeof(_lib.Weights) == 18 * 48 + 4 * 8 + 4 * 8 + 2 * 8 + 2 * 4


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_compute_fails_loudly_without_gpu():
    lib = _lib.load()
    ctx = C.c_void_p()
    rc = lib.b200al_ctx_create(C.byref(ctx), 0, 2, 1, 256, 256, 8, 0)
    assert rc == -2 and not ctx.value
    assert b"no CPU fallback" in lib.b200al_last_error(None)
    with pytest.raises(RuntimeError):
        b200al.Engine("cpu", 2)
    with pytest.raises(RuntimeError):
        b200al.Engine("cuda:0", 2)


def test_ctx_create_rejects_unsupported_configs():
    lib = _lib.load()
    ctx = C.c_void_p()
    assert lib.b200al_ctx_create(C.byref(ctx), 0, 2, 3, 256, 256, 8, 0) == -4      # RGB input
    assert lib.b200al_ctx_create(C.byref(ctx), 0, 1, 1, 256, 256, 8, 0) == -4      # K < 2
    assert lib.b200al_ctx_create(C.byref(ctx), 0, 2, 1, 200, 200, 8, 0) == -4      # untileable size
    assert lib.b200al_ctx_create(C.byref(ctx), 0, 2, 1, 256, 256, 0, 0) == -1
    assert lib.b200al_ctx_create(None, 0, 2, 1, 256, 256, 8, 0) == -1


def test_validate_state_dict():
    from oracle.unet_ref import make_state_dict
    sd = make_state_dict(0, 3)
    assert b200al.validate_state_dict(sd) == 3
    bad = dict(sd)
    del bad["encoder.levels.0.0.all.2.running_mean"]
    with pytest.raises(NotImplementedError, match="batch"):
        b200al.validate_state_dict(bad)
    bad = dict(sd)
    bad["encoder.levels.0.0.all.0.weight"] = torch.zeros(32, 3, 3, 3)
    with pytest.raises(NotImplementedError, match="grayscale"):
        b200al.validate_state_dict(bad)
    bad = dict(sd)
    bad["encoder.levels.1.0.downsample_skip.0.weight"] = torch.zeros(1)
    with pytest.raises(NotImplementedError, match="residual"):
        b200al.validate_state_dict(bad)
    bad = dict(sd)
    bad["encoder.levels.2.0.all.0.weight"] = torch.zeros(128, 64, 3, 3, 3)
    with pytest.raises(NotImplementedError, match="3D"):
        b200al.validate_state_dict(bad)


def test_selector_ctor_contract():
    s = b200al.B200EntropySelector(batch_size=12, num_workers=1, pin_memory=False)
    assert isinstance(s, b200al.ActiveSelector) and s.mode == "entropy"
    assert b200al.B200MCDropoutEntropySelector(8, 0).mc_repeats == 8
    assert b200al.B200CoresetSelector(8, 0, coreset_criteria="mean").coreset_criteria == "mean"
    with pytest.raises(RuntimeError, match="undefined"):   # coreset_selector.py:38-40
        b200al.B200CoresetSelector(8, 0, coreset_criteria="median")
    assert str(b200al.B200CoresetSelector(8, 0, feature_path="/tmp/x", coreset_fusion="cat").feature_path) == "/tmp/x"
    with pytest.raises(NotImplementedError):
        b200al.B200CoresetSelector(8, 0, coreset_fusion="mul")
    with pytest.raises(ValueError):
        b200al.B200KMeanSelector(8, 0, loaded_feature_only=True)
    assert b200al.B200DiceUncertaintySelector(8, 0).smooth == 1e-5
    for name, ctor in b200al.SELECTORS.items():
        assert isinstance(ctor(batch_size=4, num_workers=0, pin_memory=False), b200al.ActiveSelector), name


def test_cold_start_matches_reference_rng_sequence():
    """labeled_size == 0: one torch.rand(pool_size) draw on the global CPU
    generator, sort descending, first n names (entropy_selector.py:64-71)."""
    from oracle.synth import make_active
    ad = make_active(torch.zeros(20, 4, 4, dtype=torch.uint8), 0)
    torch.manual_seed(1337)
    got = b200al.B200EntropySelector(4, 0, False).select_next_batch(ad, 5, None, torch.device("cpu"))
    torch.manual_seed(1337)
    scores = torch.rand(20)
    want = [ad.pool_dataset.image_idx[i] for i in torch.sort(scores, descending=True)[1][:5]]
    assert got == want
    torch.manual_seed(1337)
    assert b200al.B200CoresetSelector(4, 0, False).select_next_batch(ad, 5, None, torch.device("cpu")) == want


def test_shard_range_partitions_pool():
    for n in (0, 1, 7, 100, 100_000):
        for w in (1, 2, 3, 8):
            r = [D.shard_range(n, i, w) for i in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1


def test_merge_topk_is_exact_with_positional_tiebreak():
    s = torch.tensor([0.5, 0.9, 0.5, 0.9, 0.1])
    p = torch.tensor([40, 7, 3, 2, 1])
    vs, ps = D.merge_topk(s, p, 3)
    assert ps.tolist() == [2, 7, 3] and vs.tolist() == [s[3].item(), s[1].item(), 0.5]
    rng = np.random.default_rng(0)
    sc = torch.from_numpy(rng.integers(0, 50, 1000).astype(np.float32))
    pos = torch.from_numpy(rng.permutation(1000))
    vs, ps = D.merge_topk(sc, pos, 100)
    from oracle.select_ref import rank_desc_stable
    full = np.empty(1000, np.float32)
    full[pos.numpy()] = sc.numpy()
    assert ps.tolist() == rank_desc_stable(full, 100).tolist()


def test_install_into_al_trainer_shim():
    class Cfg:
        active_selector_name = "b200-margin"
        batch_size, num_workers, pin_memory = 12, 1, False

    class Trainer:
        config = Cfg()

        def _setup_active_selector(self):
            raise ValueError("ActiveSelector not found")

    b200al.install_into_al_trainer(Trainer)
    t = Trainer()
    t._setup_active_selector()
    assert isinstance(t.active_selector, b200al.B200MarginSelector) and t.active_selector.batch_size == 12
    t.config.active_selector_name = "b200-coreset-l2"   # al_trainer.py:823-833 forwards the coreset options
    t.config.coreset_criteria = "mean"
    t._setup_active_selector()
    assert isinstance(t.active_selector, b200al.B200CoresetSelector)
    assert (t.active_selector.metric, t.active_selector.coreset_criteria) == ("l2", "mean")
    # al_trainer.py:845-870 also hands feature_path / loaded_feature_weight / loaded_feature_only to KMeanSelector
    t.config.active_selector_name = "b200-kmean-cosine"
    t.config.feature_path, t.config.loaded_feature_weight, t.config.loaded_feature_only = "/tmp/feats", 0.3, True
    t.config.kmean_sharp_factor, t.config.kmean_softmax = 2.0, True
    t._setup_active_selector()
    k = t.active_selector
    assert isinstance(k, b200al.B200KMeanSelector) and str(k.feature_path) == "/tmp/feats"
    assert (k.loaded_feature_weight, k.loaded_feature_only, k.sharp_factor, k.softmax) == (0.3, True, 2.0, True)
    t.config.feature_path = None   # loaded_feature_only without any loaded features: refused, not silently ignored
    with pytest.raises(ValueError):
        t._setup_active_selector()
    t.config.active_selector_name = "nope"
    with pytest.raises(ValueError):
        t._setup_active_selector()


def test_loaded_feature_files_and_cat_scale(tmp_path):
    """feature_path rows from .npy / .npz files (the .h5 stand-in where h5py is absent) and the float32 scale of
    the "cat" / mix paths (numpy < 2 rule of the reference, pyproject.toml:21)."""
    from b200al import selectors as SM
    from oracle.select_ref import cat_scale
    rows = {"a": np.arange(5, dtype=np.float32), "b": np.ones(5, np.float64)}
    np.save(tmp_path / "a.npy", rows["a"])
    np.savez(tmp_path / "b.npz", feature=rows["b"])
    got = SM._loaded_feature_rows(["b", "a"], None, tmp_path)
    assert got.dtype == np.float32 and got.tolist() == [rows["b"].tolist(), rows["a"].tolist()]
    with pytest.raises(FileNotFoundError):
        SM._loaded_feature_rows(["c"], None, tmp_path)
    for d1, d2, w in ((512, 200, 0.4), (512, 512, 1.0), (512, 700, 0.05)):
        s = SM._cat_scale(d1, d2, w)
        assert isinstance(s, np.float32) and s == cat_scale(d1, d2, w, "legacy") == np.float32(np.sqrt(d1 / d2 * w))
    x = torch.zeros(3, 712)
    assert SM._pad_to_kernel_width(x).shape == (3, 1024) and SM._pad_to_kernel_width(torch.zeros(2, 1212)).shape == (2, 1536)
    with pytest.raises(NotImplementedError):
        SM._pad_to_kernel_width(torch.zeros(1, 4096))
    with pytest.raises(RuntimeError, match="non-finite"):
        SM._require_finite(torch.tensor([1.0, float("nan")]), "embeddings")
    with pytest.raises(RuntimeError):
        SM._check_picks([3, -1], "k-center")


def test_pool_source_serves_cached_rows_in_pool_order():
    """Host side of the device pool cache (ragged packing, name -> row mapping, chunking, cache reuse); the kernel
    is replaced by a stand-in that unpacks the same buffers and runs the oracle pipeline."""
    from b200al.selectors import PoolSource
    from oracle.prep_ref import prep_image, synth_raw_u8
    from oracle.synth import MirrorExtendable, SynthRawBase
    raws = [synth_raw_u8(i) for i in range(6)]
    base = SynthRawBase(raws, image_size=48, normalize=True)
    pool = MirrorExtendable(base, base.names[2:])  # rows 2..5 are the pool

    class StandIn:
        def prep(self, raw, offsets, hw, max_h, max_w, normalize):
            assert max_h == max(r.shape[0] for r in raws) and max_w == max(r.shape[1] for r in raws)
            imgs = []
            for o, (h, w) in zip(offsets.tolist(), hw.tolist()):
                imgs.append(torch.from_numpy(prep_image(raw[o:o + h * w].reshape(h, w).numpy(), (48, 48), False)))
            x = torch.stack(imgs)
            norm = torch.stack([x.mean((1, 2)), x.std((1, 2))], 1) if normalize else None
            return x, norm

    src = PoolSource(pool, "cpu", 4, 0, False)
    assert src.cache is not None and src.image_hw() == (48, 48) and src.cache.n == 6
    got = list(src.chunks([3, 0, 1], 2, StandIn()))
    assert [g[0] for g in got] == [0, 2] and [g[1].shape[0] for g in got] == [2, 1]
    x = torch.cat([g[1] for g in got])
    norm = torch.cat([g[2] for g in got])
    for k, pos in enumerate([3, 0, 1]):
        want = pool.get_sample(pos)["image"][0]  # oracle pipeline incl. z-score
        z = (x[k] - norm[k, 0]) / norm[k, 1]
        assert (z - want).abs().max() < 1e-5
    assert PoolSource(pool, "cpu", 4, 0, False).cache is src.cache          # built once per run
    assert PoolSource(pool, "cpu", 4, 0, False, pool_cache="off").cache is None
    base.transform = lambda d: d                                             # random transforms: never cached
    base._b200_pool_cache = None
    assert PoolSource(pool, "cpu", 4, 0, False).cache is None


def test_conv_flops_match_survey():
    assert abs(b200al.conv_flops_per_sample() / 1e9 - (24.071 - 0.038 - 0.008)) < 0.02   # minus first conv and head
    assert abs(b200al.conv_flops_per_sample(encoder_only=True) / 1e9 - (8.493 - 0.038)) < 0.01


def test_bench_reference_arm_prints_exactly_one_json_line():
    """bench.py's contract: stdout is ONE JSON line (libraries that write to fd 1 are pointed at stderr).  The
    reference arm runs on the CPU, so it can be checked here; the keys are the ones the driver reads."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0", "--cpu-sample", "1"], capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "images/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"]
