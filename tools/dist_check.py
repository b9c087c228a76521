"""Selections must not depend on how many GPUs share the pool.  Run once plainly and once under torchrun, then compare:

    python tools/dist_check.py > one.json
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py > two.json
    cmp one.json <(grep -v '^NCCL' two.json)

Every drop-in selector picks a budget from a 300-image synthetic pool (K=3 golden model); rank 0 prints the names.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "medical-image-analysis_b200")]

import numpy as np
import torch

import b200al
from oracle.synth import make_active, synth_pool_u8
from oracle.unet_ref import OracleUNet, golden_state_dict


def main():
    if "RANK" in os.environ:
        torch.distributed.init_process_group("nccl")
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    rank = int(os.environ.get("RANK", 0))
    dev = torch.device("cuda", torch.cuda.current_device())
    g = np.load(os.path.join(ROOT, "tests", "golden", "model.npz"))
    model = OracleUNet(golden_state_dict(3, g))
    images = synth_pool_u8(20 + 301)  # odd pool size: uneven shards
    kw = dict(batch_size=8, num_workers=0, pin_memory=False, chunk_samples=64)
    out = {}
    for name in sorted(b200al.SELECTORS):
        extra = dict(mc_repeats=4, dropout_prob=0.1) if name == "b200-mc-entropy" else {}
        sel = b200al.SELECTORS[name](**kw, **extra)
        np.random.seed(31)
        torch.manual_seed(31)
        out[name] = sel.select_next_batch(make_active(images, 20), 12, model, dev)
    # k-center variants: under torchrun the default path is the sharded persistent kernel (one candidate per rank and
    # step over NVLink); `shard_kcenter=False` gathers the rows and runs replicated - all must equal the 1-GPU picks
    for metric in ("l2", "cosine"):
        for crit in ("min", "mean"):
            for shard in (True, False):
                sel = b200al.B200CoresetSelector(metric=metric, coreset_criteria=crit, shard_kcenter=shard, **kw)
                out[f"coreset/{metric}/{crit}/{'sharded' if shard else 'replicated'}"] = \
                    sel.select_next_batch(make_active(images, 20), 40, model, dev)
    for k in [k for k in out if k.endswith("/sharded")]:
        assert out[k] == out[k.replace("/sharded", "/replicated")], k
    if rank == 0:
        print(json.dumps(out, indent=0, sort_keys=True))
    if "RANK" in os.environ:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
