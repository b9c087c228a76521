"""Time of the k-means++ seeding kernels alone (b200al_kmeans_pp) at the BASELINE pool sizes."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]
import numpy as np
import torch

import b200al

eng = b200al.Engine("cuda:0", 2, max_samples=1)
g = torch.Generator(device="cuda:0").manual_seed(1)
for n, k in ((10_000, 100), (100_000, 500)):
    X = eng.standardize_rows(torch.randn((n, 512), device="cuda:0", generator=g))
    w = torch.rand(n, generator=torch.Generator().manual_seed(2))
    w /= w.sum()
    trials = 2 + int(np.log(k))
    u = torch.from_numpy(np.random.RandomState(0).uniform(size=(k - 1, trials)))
    eng.kmeans_pp(X, w, 5, u, k, trials)
    torch.cuda.synchronize()
    t = time.perf_counter()
    p = eng.kmeans_pp(X, w, 5, u, k, trials)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print("kmeans_pp n=%d k=%d trials=%d: %.3f s (%.1f us/centre), distinct %d" % (n, k, trials, dt, dt / k * 1e6, len(set(p.tolist()))))
