"""b200al_topk_static: rank & cut through the C ABI with NO allocation (scratch in a file-scope static __device__
array of the library, include/b200al.h).  Checked against the existing b200al_topk on two calls in a row; collected
last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _static_topk(lib, scores: torch.Tensor, k: int, base_pos: int):
    """Outputs are the caller's (allocated here, BEFORE the call); the entry point itself allocates nothing."""
    n = scores.numel()
    kk = min(k, n)
    vals = torch.empty(kk, dtype=torch.float32, device=scores.device)
    pos = torch.empty(kk, dtype=torch.int64, device=scores.device)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(scores.device)[0]
    rc = lib.b200al_topk_static(scores.device.index, C.c_void_p(scores.data_ptr()), n, k, base_pos,
                                C.c_void_p(vals.data_ptr()), C.c_void_p(pos.data_ptr()),
                                C.c_void_p(torch.cuda.current_stream(scores.device).cuda_stream))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(scores.device)[0]
    return rc, vals, pos, free0 - free1


def test_topk_static_matches_topk_on_two_calls_in_a_row():
    import b200al
    from b200al import _lib
    lib = _lib.load()
    eng = b200al.Engine(DEV, 2, max_samples=1)
    try:
        rng = np.random.default_rng(17)
        # call 1: large, heavy ties; call 2: smaller and different (stale keys / candidates / state of call 1 remain
        # in the shared array); call 3: the maximum size
        cases = [(rng.integers(0, 40, 200_003).astype(np.float32), 5000, 1000),
                 (rng.standard_normal(10_007).astype(np.float32), 100, 0),
                 (rng.standard_normal(_lib.TOPK_STATIC_MAX_N).astype(np.float32), 500, 7)]
        for i, (sc, k, base) in enumerate(cases):
            d = torch.from_numpy(sc).to(DEV)
            want_v, want_p = eng.topk(d, k, base_pos=base)          # existing entry point (b200al_topk)
            rc, vals, pos, grew = _static_topk(lib, d, k, base)
            assert rc == 0, lib.b200al_last_error(None)
            assert torch.equal(pos, want_p) and torch.equal(vals, want_v), f"call {i}"   # exact, like the top-k tests
            if i > 0:   # (the first launch of a kernel may load its module)
                assert grew <= 0, f"device memory in use grew by {grew} bytes across a b200al_topk_static call"
        # k > n clips, k == 0 and n == 0 are no-ops - same contract as b200al_topk
        d = torch.from_numpy(cases[1][0][:5].copy()).to(DEV)
        rc, vals, pos, _ = _static_topk(lib, d, 9, 3)
        wv, wp = eng.topk(d, 9, base_pos=3)
        assert rc == 0 and torch.equal(pos, wp) and torch.equal(vals, wv)
        assert lib.b200al_topk_static(0, None, 0, 5, 0, None, None, None) == 0
        # beyond the static scratch: rejected with a clear error, nothing launched
        big = torch.zeros(_lib.TOPK_STATIC_MAX_N + 1, device=DEV)
        rc, _, _, _ = _static_topk(lib, big, 10, 0)
        assert rc == -1
        msg = lib.b200al_last_error(None).decode()
        assert "B200AL_TOPK_STATIC_MAX_N" in msg and str(_lib.TOPK_STATIC_MAX_N) in msg
    finally:
        eng.close()
