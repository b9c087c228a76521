"""World-size-2 gloo run of the sharded selection logic (CPU; scores come from
a stand-in so no GPU is needed): shard -> local top-k -> all_gather -> merge
must equal the single-process ranking, and gathered rows keep pool order."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200al import dist as D


def _worker(rank, world, port, n, k, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        scores = torch.from_numpy(rng.integers(0, 40, n).astype(np.float32))  # many ties
        feats = torch.from_numpy(rng.standard_normal((n, 8)).astype(np.float32))
        lo, hi = D.shard_range(n, rank, world)
        local = scores[lo:hi]
        o = torch.argsort(local, descending=True, stable=True)[:k]
        vals, pos = D.global_topk(local[o], o + lo, k)
        rows = D.gather_rows(feats[lo:hi])
        if rank == 0:
            torch.save({"pos": pos, "vals": vals, "rows_equal": torch.equal(rows, feats)}, out)
    finally:
        dist.destroy_process_group()


def test_sharded_topk_and_gather_world2(tmp_path):
    n, k = 1001, 37
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, 29517, n, k, out), nprocs=2, join=True)
    got = torch.load(out)
    rng = np.random.default_rng(5)
    scores = rng.integers(0, 40, n).astype(np.float32)
    from oracle.select_ref import rank_desc_stable
    assert got["pos"].tolist() == rank_desc_stable(scores, k).tolist()
    assert got["rows_equal"]


def _bcast_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(100 + rank)  # every rank starts from DIFFERENT weights
        sd = {"a.weight": torch.randn(3, 4, generator=g), "a.bias": torch.randn(3, generator=g),
              "bn.num_batches_tracked": torch.tensor(7 + rank)}
        got = D.broadcast_state_dict(sd, "cpu")
        torch.save({k: v.clone() for k, v in got.items()}, f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


def test_state_dict_broadcast_world2(tmp_path):
    out = str(tmp_path / "sd")
    mp.spawn(_bcast_worker, args=(2, 29519, out), nprocs=2, join=True)
    r0, r1 = torch.load(f"{out}.0"), torch.load(f"{out}.1")
    g = torch.Generator().manual_seed(100)
    want_w, want_b = torch.randn(3, 4, generator=g), torch.randn(3, generator=g)
    for r in (r0, r1):
        assert torch.equal(r["a.weight"], want_w) and torch.equal(r["a.bias"], want_b)
    assert int(r0["bn.num_batches_tracked"]) == 7 and int(r1["bn.num_batches_tracked"]) == 8  # non-float: local


def test_single_process_passthrough():
    assert D.world() == (0, 1)
    t = torch.arange(5)
    assert D.all_gather_varlen(t) is t
    sd = {"w": torch.ones(2)}
    assert D.broadcast_state_dict(sd, "cpu") is sd


def _draws_worker(rank, world, port, out):
    """Ranks whose RNG states DIFFER (rank-offset seeds) must still agree on the cold-start picks and on the k-means++
    draws: rank 0's values are broadcast (selectors._cold_start, dist.broadcast_draws); every rank still consumes its
    own stream, like the single-process reference does."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import b200al
        from oracle.synth import make_active
        torch.manual_seed(1000 + rank)
        np.random.seed(2000 + rank)
        ad = make_active(torch.zeros(30, 4, 4, dtype=torch.uint8), 0)
        cold = b200al.B200EntropySelector(4, 0, False).select_next_batch(ad, 6, None, torch.device("cpu"))
        after_cold = float(torch.rand(1))                     # own stream consumed: differs per rank
        first = int(np.random.mtrand._rand.choice(30))
        uni = np.random.mtrand._rand.uniform(size=(5, 3))
        f2, u2 = D.broadcast_draws(first, uni)
        ranges = [D.shard_range(101, r, world) for r in range(world)]
        torch.save({"cold": cold, "after_cold": after_cold, "first": f2, "uni": u2, "own_first": first,
                    "bcast": D.broadcast_cpu(torch.arange(4) + 10 * rank).tolist(), "ranges": ranges}, f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


def test_cold_start_and_kmeanspp_draws_agree_across_ranks_world2(tmp_path):
    out = str(tmp_path / "draws")
    mp.spawn(_draws_worker, args=(2, 29521, out), nprocs=2, join=True)
    r0, r1 = torch.load(f"{out}.0", weights_only=False), torch.load(f"{out}.1", weights_only=False)
    torch.manual_seed(1000)
    want = [f"{i:06d}" for i in torch.sort(torch.rand(30), descending=True)[1][:6].tolist()]
    assert r0["cold"] == r1["cold"] == want                   # rank 0's draw, the reference's RNG call sequence
    assert r0["after_cold"] != r1["after_cold"]
    assert r0["first"] == r1["first"] == r0["own_first"] and np.array_equal(r0["uni"], r1["uni"])
    assert r0["bcast"] == r1["bcast"] == [0, 1, 2, 3]
    assert r0["ranges"] == [(0, 50), (50, 101)]
