"""How well-conditioned is the reference's k-center selection?  (container only: needs the reference embeddings that
oracle/make_golden_configs.py leaves under $B200AL_GOLDEN_TMP.)

The reference's own config-3 run at N = 20 500 has per-step top-2 gaps down to 4.8e-7 - one fp32 ulp of the distance -
so its pick ORDER is decided by the last bit of a BLAS result.  This script perturbs the reference embeddings by
increasing amounts of noise and re-runs the streaming k-center oracle: where does the order first change, and how much
of the selected SET survives?  Record: profiles/r02_kcenter_conditioning.json.

  python tools/kcenter_conditioning.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.make_golden_configs import BUDGET3, N_LABELED3, TMP, _stream_l2  # noqa: E402


def main():
    emb = np.load(os.path.join(TMP, "fwd20k_emb.npy"))
    base, gaps = _stream_l2(emb, N_LABELED3, BUDGET3)
    rng = np.random.default_rng(0)
    am = float(np.abs(emb).max())
    up = rng.random(emb.shape) < 0.5
    cases = {
        "1 fp32 ulp, random direction": np.where(up, np.nextafter(emb, np.float32(np.inf)), np.nextafter(emb, np.float32(-np.inf))),
        "relative 1e-6": emb * (1 + 1e-6 * rng.standard_normal(emb.shape)),
        "relative 1e-5": emb * (1 + 1e-5 * rng.standard_normal(emb.shape)),
        "relative 1e-4": emb * (1 + 1e-4 * rng.standard_normal(emb.shape)),
        "absolute 1e-4 x absmax": emb + 1e-4 * am * rng.standard_normal(emb.shape),
        "relative 1e-3": emb * (1 + 1e-3 * rng.standard_normal(emb.shape)),
    }
    out = {"rows": int(emb.shape[0]), "labeled": N_LABELED3, "budget": BUDGET3, "golden_min_top2_gap": float(min(gaps)),
           "golden_steps_with_gap_below_1e-5": int((np.array(gaps) < 1e-5).sum()), "perturbations": {}}
    for name, e in cases.items():
        p, _ = _stream_l2(e.astype(np.float32), N_LABELED3, BUDGET3)
        first = next((i for i, (a, b) in enumerate(zip(base, p)) if a != b), None)
        out["perturbations"][name] = {"first_order_divergence_step": first, "set_overlap_of_500": len(set(base) & set(p))}
        print(name, out["perturbations"][name], flush=True)
    json.dump(out, open(os.path.join(ROOT, "profiles", "r02_kcenter_conditioning.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
