"""Pool sharding across ranks (one process per GPU, torch.distributed).

The reference is single-process; sharding is new.  Images are
independent on the scoring path (entropy_selector.py:38-52 has no cross-image
term, BatchNorm is in eval mode), so the pool is cut into contiguous position
ranges with NO data-path collective; only per-shard top-k candidates
(score fp32, position int64) or embedding rows are exchanged:

  uncertainty: all_gather of each rank's top-min(k, n_r) (score, pos) pairs,
               every rank merges identically (score desc, position asc).
  coreset:     all_gather of embedding rows [n_r, 512]; k-center then runs
               replicated on every rank (500 greedy steps are latency-bound:
               a per-step collective would cost more than the whole loop).

Works with the gloo backend on CPU tensors (tests) and nccl on CUDA tensors.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous [lo, hi) of rank `rank`: floor(r*n/W) .. floor((r+1)*n/W)."""
    return (rank * n) // world_size, ((rank + 1) * n) // world_size


def merge_topk(scores: torch.Tensor, pos: torch.Tensor, k: int):
    """Exact merge of candidate lists: descending score, ascending position."""
    if scores.numel() == 0:
        return scores, pos
    # two stable sorts: by position, then by score descending
    o = torch.argsort(pos, stable=True)
    scores, pos = scores[o], pos[o]
    o = torch.argsort(scores, descending=True, stable=True)
    return scores[o][:k], pos[o][:k]


def all_gather_varlen(t: torch.Tensor, group=None) -> torch.Tensor:
    """Concatenate per-rank tensors of differing first dimension (rank order)."""
    rank, ws = world()
    if ws == 1:
        return t
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    sizes = [torch.zeros_like(n) for _ in range(ws)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    bufs = [torch.empty_like(pad) for _ in range(ws)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:s] for b, s in zip(bufs, sizes)], 0)


def global_topk(local_scores: torch.Tensor, local_pos: torch.Tensor, k: int, group=None):
    """local_* = this rank's own top-min(k, n_r) candidates with GLOBAL pool
    positions.  Returns the global top-k, identical on every rank."""
    s = all_gather_varlen(local_scores, group)
    p = all_gather_varlen(local_pos, group)
    return merge_topk(s, p, k)


def gather_rows(local_rows: torch.Tensor, group=None) -> torch.Tensor:
    """All ranks' rows in rank (== pool position) order."""
    return all_gather_varlen(local_rows, group)


def broadcast_state_dict(sd: dict, device, src: int = 0, group=None) -> dict:
    """Rank `src`'s floating-point tensors of `sd` on every rank (one flat fp32 broadcast), so that all ranks score
    their shard with the same weights even if their trainers drifted; non-float entries (num_batches_tracked) are
    kept local.  The reference reloads / rebuilds the model before each selection (al_trainer.py:1048-1083)."""
    rank, ws = world()
    if ws == 1:
        return sd
    names = [k for k, v in sd.items() if torch.is_tensor(v) and v.is_floating_point()]
    flat = torch.cat([sd[k].detach().to(device=device, dtype=torch.float32).reshape(-1) for k in names])
    dist.broadcast(flat, src=src, group=group)
    out, off = dict(sd), 0
    for k in names:
        n = sd[k].numel()
        out[k] = flat[off:off + n].view(sd[k].shape)
        off += n
    return out


def broadcast_cpu(t: torch.Tensor, src: int = 0, group=None) -> torch.Tensor:
    """Rank `src`'s copy of a small CPU tensor on every rank (same shape everywhere); identity when not distributed.
    NCCL groups move it through the current CUDA device."""
    rank, ws = world()
    if ws == 1:
        return t
    if dist.get_backend(group) == "nccl":
        d = t.to(torch.device("cuda", torch.cuda.current_device()))
        dist.broadcast(d, src=src, group=group)
        return d.cpu()
    t = t.clone()
    dist.broadcast(t, src=src, group=group)
    return t


def broadcast_draws(first: int, uniforms, device=None, src: int = 0, group=None):
    """k-means++ random draws (first centre, uniforms [(k-1), trials] float64 numpy): rank `src`'s values on every
    rank.  Each rank still consumes its own numpy stream like the single-process reference; only the values agree."""
    import numpy as np
    rank, ws = world()
    if ws == 1:
        return first, uniforms
    flat = torch.from_numpy(np.concatenate([[float(first)], np.asarray(uniforms, np.float64).reshape(-1)]))
    flat = broadcast_cpu(flat, src, group)
    return int(flat[0].item()), flat[1:].numpy().reshape(np.asarray(uniforms).shape)


# ---------------------------------------------------------------------------------------------------------------
# Peer-accessible buffers for the sharded k-center kernel (b200al_kcenter_run_sharded): one symmetric byte buffer per
# rank, [mailbox | embedding shard rows], mapped into every process of the group by torch's symmetric memory
# (NVLink P2P / fabric handles).  torch only allocates and exchanges the handles; the kernel does the exchange itself.
# ---------------------------------------------------------------------------------------------------------------
_PEER_BUFFERS: dict = {}


class PeerBuffers:
    MAIL_ALIGN = 256
    MAX_EPOCH = 30000

    def __init__(self, device, rows_cap: int, width: int, mail_bytes: int, group=None):
        import torch.distributed._symmetric_memory as symm
        rank, ws = world()
        self.device, self.rows_cap, self.width = torch.device(device), int(rows_cap), int(width)
        self.mail_bytes = (int(mail_bytes) + self.MAIL_ALIGN - 1) // self.MAIL_ALIGN * self.MAIL_ALIGN
        nbytes = self.mail_bytes + self.rows_cap * self.width * 4
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.buf.zero_()                      # mailbox tags start at 0 ("never written")
        torch.cuda.synchronize(self.device)
        self.handle = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.ptrs = [int(x) for x in self.handle.buffer_ptrs]
        if len(self.ptrs) != ws or self.ptrs[rank] != self.buf.data_ptr():
            raise RuntimeError("symmetric memory rendezvous returned unexpected peer pointers")
        self.epoch = 0
        dist.barrier(group=group)             # every rank's zeroing has landed before anybody publishes

    def rows(self, n: int) -> torch.Tensor:
        """float32 [n, width] view of this rank's shard area."""
        if n > self.rows_cap:
            raise ValueError("shard larger than the peer buffer")
        return self.buf[self.mail_bytes:self.mail_bytes + n * self.width * 4].view(torch.float32).view(n, self.width)

    def next_epoch(self, group=None) -> int:
        self.epoch += 1
        if self.epoch > self.MAX_EPOCH:       # tags would wrap: start over with clean mailboxes
            torch.cuda.synchronize(self.device)
            dist.barrier(group=group)
            self.buf[:self.mail_bytes].zero_()
            torch.cuda.synchronize(self.device)
            dist.barrier(group=group)
            self.epoch = 1
        return self.epoch

    def mail_ptrs(self):
        return list(self.ptrs)

    def row_ptrs(self):
        return [p + self.mail_bytes for p in self.ptrs]


def peer_buffers(device, rows_cap: int, width: int, mail_bytes: int, group=None) -> PeerBuffers:
    """Cached per (device, capacity, width): the rendezvous is a collective and the mapping stays valid."""
    dev = torch.device(device)
    key = (dev.index, int(width))
    pb = _PEER_BUFFERS.get(key)
    if pb is None or pb.rows_cap < rows_cap:
        # capacities must agree across ranks: they are derived from the pool size and the world size only
        pb = _PEER_BUFFERS[key] = PeerBuffers(dev, rows_cap, width, mail_bytes, group)
    return pb
