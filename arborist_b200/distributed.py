"""Tree-set sharding across GPUs.

Candidate trees are independent (A.py:1089-1126 keeps no cross-tree state except
the running argmax), so each rank fits trees ``rank, rank + W, rank + 2W, ...``
on its own GPU against a replicated read-count matrix, and the only exchange is
one all-gather of the per-tree ELBOs (and iteration counts).  ``torch.distributed``
is plumbing only: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import numpy as np

__all__ = ["world", "shard_indices", "gather_elbos", "fetch_fit", "gather_vector"]


def world():
    """(rank, world_size) of the initialised default process group, else (0, 1)."""
    try:
        import torch.distributed as dist
    except ImportError:          # pragma: no cover
        return 0, 1
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_indices(n_trees, rank=None, world_size=None):
    """Global tree indices owned by this rank (interleaved: iteration counts vary 3..25)."""
    if rank is None or world_size is None:
        rank, world_size = world()
    return np.arange(rank, n_trees, world_size, dtype=np.int64)


def gather_vector(local, shard, n_total, dtype=np.float64):
    """All-gather a per-tree vector sharded with :func:`shard_indices`; returns the full vector."""
    rank, ws = world()
    local = np.asarray(local, dtype=dtype)
    if ws == 1:
        out = np.empty(n_total, dtype=dtype)
        out[shard] = local
        return out
    import torch
    import torch.distributed as dist

    per = (n_total + ws - 1) // ws               # equal-sized buffers for all_gather
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    buf = torch.zeros(per, dtype=torch.from_numpy(np.zeros(1, dtype)).dtype, device=dev)
    buf[: local.shape[0]] = torch.from_numpy(local).to(dev)
    parts = [torch.empty_like(buf) for _ in range(ws)]
    dist.all_gather(parts, buf)
    out = np.empty(n_total, dtype=dtype)
    for r in range(ws):
        idx = shard_indices(n_total, r, ws)
        out[idx] = parts[r][: idx.shape[0]].cpu().numpy()
    return out


def gather_elbos(elbo_local, shard, n_total):
    """Full ELBO vector + owning rank of every tree."""
    _, ws = world()
    owner = np.arange(n_total, dtype=np.int64) % ws
    return gather_vector(elbo_local, shard, n_total), owner


def fetch_fit(engine, idx, local_of, owner):
    """Posteriors of global tree ``idx``: local trees come from this rank's GPU; remote ones
    are broadcast by their owner (collective: every rank must ask for the same tree)."""
    rank, ws = world()
    if ws == 1 or int(owner[idx]) == rank:
        qz, qy = engine.get_fit(local_of[idx])
        if ws == 1:
            return qz, qy
    else:
        n, m, _, _ = engine.dims()
        qz = np.empty((n, engine.n_clones))
        qy = np.empty((m, engine.n_clones - 1))
    import torch
    import torch.distributed as dist

    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    tz, ty = torch.from_numpy(qz).to(dev), torch.from_numpy(qy).to(dev)
    dist.broadcast(tz, int(owner[idx]))
    dist.broadcast(ty, int(owner[idx]))
    return tz.cpu().numpy(), ty.cpu().numpy()
