"""Tree-set sharding across GPUs.

Candidate trees are independent (A.py:1089-1126 keeps no cross-tree state except
the running argmax), so each rank fits trees ``rank, rank + W, rank + 2W, ...``
on its own GPU against a replicated read-count matrix, and the only exchange is
one all-gather of the per-tree ELBOs (and iteration counts).  Fetching the posteriors of a
tree that ANOTHER rank fitted (``TreeFit.q_z`` of a remote tree, ``--pickle`` of all fits) is a
broadcast from its owner, i.e. a collective: every rank has to ask for the same remote tree in
the same order.  A single process that drives all GPUs itself (``arborist(..., devices=...)``,
``arb_fit_trees_multi``) needs none of this.  ``torch.distributed``
is plumbing only: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import numpy as np

__all__ = ["world", "shard_indices", "gather_elbos", "broadcast_fit", "gather_vector", "all_agree", "local_device"]


def world():
    """(rank, world_size) of the initialised default process group, else (0, 1)."""
    import sys

    if "torch" not in sys.modules:          # nobody can have initialised a process group: do not pay the import
        return 0, 1
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


def local_device(device=None):
    """CUDA ordinal this rank works on: ``device`` when given, else ``LOCAL_RANK`` modulo the number of visible GPUs
    -- the rule ``Engine(device=None)`` applies, so collectives and kernels agree before any engine exists."""
    if device is not None:
        return int(device)
    import os

    import torch

    return int(os.environ.get("LOCAL_RANK", "0")) % max(1, torch.cuda.device_count())


def _comm_device(device=None):
    """Where collective buffers live: this rank's GPU under NCCL (NOT torch's current device, which is 0 on every
    rank unless the caller ran ``torch.cuda.set_device``), the host under gloo."""
    import torch
    import torch.distributed as dist

    if dist.get_backend() != "nccl":
        return torch.device("cpu")
    return torch.device("cuda", local_device(device))


def all_agree(ok, device=None):
    """True on every rank iff ``ok`` is true on every rank (so that all ranks raise together).  ``device``: the
    caller's explicit CUDA ordinal, if it has one (see :func:`local_device`)."""
    _, ws = world()
    if ws == 1:
        return bool(ok)
    import torch
    import torch.distributed as dist

    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=_comm_device(device))
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t.item()))


def gather_vector(local, shard, n_total, dtype=np.float64, device=None):
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
    dev = _comm_device(device)
    buf = torch.zeros(per, dtype=torch.from_numpy(np.zeros(1, dtype)).dtype, device=dev)
    buf[: local.shape[0]] = torch.from_numpy(local).to(dev)
    parts = [torch.empty_like(buf) for _ in range(ws)]
    dist.all_gather(parts, buf)
    out = np.empty(n_total, dtype=dtype)
    for r in range(ws):
        idx = shard_indices(n_total, r, ws)
        out[idx] = parts[r][: idx.shape[0]].cpu().numpy()
    return out


def gather_elbos(elbo_local, shard, n_total, device=None):
    """Full ELBO vector + owning rank of every tree."""
    _, ws = world()
    owner = np.arange(n_total, dtype=np.int64) % ws
    return gather_vector(elbo_local, shard, n_total, device=device), owner


def broadcast_fit(qz, qy, owner_rank, device=None):
    """Posteriors of one tree from the rank that fitted it to every rank (COLLECTIVE: every rank calls this for
    the same tree; ``qz`` / ``qy`` are the real arrays on the owner and same-shaped buffers elsewhere)."""
    import torch
    import torch.distributed as dist

    dev = _comm_device(device)
    tz, ty = torch.from_numpy(np.ascontiguousarray(qz)).to(dev), torch.from_numpy(np.ascontiguousarray(qy)).to(dev)
    dist.broadcast(tz, int(owner_rank))
    dist.broadcast(ty, int(owner_rank))
    return tz.cpu().numpy(), ty.cpu().numpy()
