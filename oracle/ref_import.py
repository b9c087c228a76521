"""Import the unmodified reference implementation (container only).

/root/reference does not exist on the GPU box; callers must check
`reference_available()` first.  Nothing here is used at GPU-test / bench time.

The reference's `src/` holds top-level packages with generic names
(`datasets`, `models`, `utils`, `transforms`, `activelearning`).  Several of
them import third-party modules at module top that are absent here and never
touched on the acquisition path (src/datasets/__init__.py:2-6,
src/activelearning/coreset_selector.py:11).  Those are stubbed in sys.modules.
"""
from __future__ import annotations

import importlib
import os
import sys
import types
from types import SimpleNamespace

REFERENCE_SRC = os.environ.get("B200AL_REFERENCE_SRC", "/root/reference/src")

_STUBS = [
    "SimpleITK", "h5py", "medpy", "medpy.metric", "medpy.metric.binary",
    "skimage", "skimage.measure", "skimage.transform", "skimage.io",
    "skimage.morphology", "skimage.segmentation", "skimage.filters",
    "monai", "gradio", "open_clip", "nnunetv2", "nrrd", "wandb", "nibabel",
]

_cache: SimpleNamespace | None = None


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "activelearning"))


class _Anything:
    """Attribute sink: any attribute is a class usable in annotations/isinstance."""

    def __init__(self, *a, **k):
        pass


def _install_stubs() -> None:
    for name in _STUBS:
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
            continue
        except Exception:
            pass
        mod = types.ModuleType(name)

        def _getattr(attr, _n=name):
            if attr.startswith("__"):
                raise AttributeError(attr)
            return _Anything

        mod.__getattr__ = _getattr  # type: ignore[attr-defined]
        mod.__path__ = []  # behave like a package
        sys.modules[name] = mod
        if "." in name:
            parent, child = name.rsplit(".", 1)
            setattr(sys.modules[parent], child, mod)


def load_reference() -> SimpleNamespace:
    """Returns the reference's selector / model / dataset classes, untouched."""
    global _cache
    if _cache is not None:
        return _cache
    if not reference_available():
        raise RuntimeError(f"reference sources not found at {REFERENCE_SRC}")
    for generic in ("datasets", "models", "utils", "transforms", "activelearning", "losses"):
        m = sys.modules.get(generic)
        if m is not None and not str(getattr(m, "__file__", "")).startswith(REFERENCE_SRC):
            raise RuntimeError(
                f"module '{generic}' already imported from {getattr(m, '__file__', '?')}; "
                "import the reference in a fresh process")
    import torch  # noqa: F401  (import the real heavy modules before stubbing)
    import torchvision  # noqa: F401
    import sklearn.metrics  # noqa: F401
    import sklearn.cluster  # noqa: F401

    _install_stubs()
    sys.path.insert(0, REFERENCE_SRC)
    try:
        al = importlib.import_module("activelearning")
        unet_mod = importlib.import_module("models.unet.unet")
        ds = importlib.import_module("datasets")
        base = importlib.import_module("datasets.basedataset")
        cs = importlib.import_module("activelearning.coreset_selector")
        norm = importlib.import_module("transforms.normalization")
        jt = importlib.import_module("transforms.joint_transform")
        dl = importlib.import_module("losses.dice_loss")
    finally:
        sys.path.remove(REFERENCE_SRC)
    _cache = SimpleNamespace(
        ActiveSelector=al.ActiveSelector,
        RandomSelector=al.RandomSelector,
        EntropySelector=al.EntropySelector,
        MarginSelector=al.MarginSelector,
        ConfidenceSelector=al.ConfidenceSelector,
        CoresetSelector=al.CoresetSelector,
        KMeanSelector=al.KMeanSelector,
        kcenter_greedy=cs.kcenter_greedy,
        UNet=unet_mod.UNet,
        ActiveDataset=ds.ActiveDataset,
        ExtendableDataset=ds.ExtendableDataset,
        BaseDataset=base.BaseDataset,
        FUGCDataset=ds.FUGCDataset,
        BUSIDataset=ds.BUSIDataset,
        ZScoreNormalize=norm.ZScoreNormalize,
        JointResize=jt.JointResize,
        DiceLoss=dl.DiceLoss,
    )
    return _cache


def build_reference_unet(num_classes_plus_bg: int, in_channels: int = 1, dropout_prob: float | None = 0.1):
    """The exact construction al_train uses (src/training/al_trainer.py:504-515)."""
    ref = load_reference()
    return ref.UNet(
        dimension=2,
        input_channels=in_channels,
        output_classes=num_classes_plus_bg,
        channels_list=[32, 64, 128, 256, 512],
        deep_supervision=False,
        ds_layer=0,
        block_type="plain",
        dropout_prob=dropout_prob,
        normalization="batch",
    )
