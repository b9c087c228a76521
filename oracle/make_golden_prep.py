"""Generate tests/golden/prep.npz by running the UNMODIFIED reference dataset pipeline here (container only; needs
/root/reference, PIL and torchvision).  Usage:  python -m oracle.make_golden_prep

Seeded ragged grayscale images (oracle.prep_ref.synth_raw_u8, indices 0..2: FUGC-sized 336x544, a larger BUSI-like
one, a small one that is UP-sampled) are written as PNGs in the FUGC directory layout, read back through the
reference's `FUGCDataset(split="train", transform=None, image_channels=1, image_size=256)` with and without
`ZScoreNormalize()` - exactly how `al_train` builds the pool dataset (al_trainer.py:552-605,623) - and the
resulting tensors are stored:
  resized    [3,256,256] fp32   dataset output without normalisation
  mean, std  [3]                torch's `image.mean()`, `image.std()` of those (normalization.py:18-19)
  zsum       [3] fp64           sum of |normalised image| (checksum of the --do-normalize output)
  zprobe     [3,16,16] fp32     normalised image sampled every 16th row / column
The raw inputs are regenerated from the seed at test time, not stored.
"""
from __future__ import annotations

import os
import tempfile

import numpy as np
import torch

from .prep_ref import synth_raw_u8
from .ref_import import load_reference

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "prep.npz")
N = 3


def write_fugc_tree(root: str, raws: list, names: list[str]) -> None:
    from PIL import Image
    for sub in ("images", "labels"):
        os.makedirs(os.path.join(root, "train", sub), exist_ok=True)
    for raw, name in zip(raws, names):
        Image.fromarray(raw, mode="L").save(os.path.join(root, "train", "images", f"{name}.png"))
        Image.fromarray(np.zeros_like(raw), mode="L").save(os.path.join(root, "train", "labels", f"{name}.png"))


def reference_samples(root: str, names: list[str], normalize: bool):
    ref = load_reference()
    ds = ref.FUGCDataset(data_path=root, split="train", normalize=ref.ZScoreNormalize() if normalize else None,
                         transform=None, image_channels=1, image_size=256)
    by_name = {ds.get_sample(i)["case_name"]: i for i in range(len(ds))}  # glob order is arbitrary
    return ds, [ds.get_sample(by_name[n])["image"][0] for n in names]


def main():
    import torchvision
    raws = [synth_raw_u8(i) for i in range(N)]
    names = [f"case{i:03d}" for i in range(N)]
    with tempfile.TemporaryDirectory() as root:
        write_fugc_tree(root, raws, names)
        _, plain = reference_samples(root, names, False)
        _, zs = reference_samples(root, names, True)
    resized = torch.stack(plain)
    np.savez(OUT,
             resized=resized.numpy(),
             mean=np.array([float(x.mean()) for x in plain], np.float32),
             std=np.array([float(x.std()) for x in plain], np.float32),
             zsum=np.array([float(z.double().abs().sum()) for z in zs], np.float64),
             zprobe=torch.stack(zs)[:, ::16, ::16].numpy(),
             sizes=np.array([r.shape for r in raws], np.int32),
             versions=np.array([f"torch={torch.__version__}", f"torchvision={torchvision.__version__}"]))
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
