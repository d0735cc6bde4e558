"""Result materialisation at scale.

The reference returns ``all_tree_fits = {tree_idx: TreeFit}`` with every tree's
``q_z`` / ``q_y`` held in host memory (A.py:1109-1120) and pickles that dict in one
go (A.py:1187-1188).  At the headline shapes that is 2.7 GB (config 3) to 236 GB
(config 4), so here

* :class:`LazyFits` is a ``dict`` whose ``TreeFit`` values are created on first
  access and whose posteriors stay on the GPU until touched;
* :class:`FitStore` maps a global tree index to the engine that fitted it; when the
  posteriors were not kept on the device it makes BATCHES of consecutive local trees
  resident again with one batched re-fit (``arb_get_fits``) and copies single trees
  out of the batch;
* :func:`write_fits_pickle` streams the dict to disk as ONE ordinary pickle
  (``pd.read_pickle`` gives ``{idx: TreeFit}`` with plain arrays and the four
  index dicts shared between the fits, like the reference's file) while holding
  one batch of posteriors at a time.
"""
from __future__ import annotations

import pickle

import numpy as np

__all__ = ["FitStore", "LazyFits", "write_fits_pickle"]


class FitStore:
    """Where the posteriors of every tree of one ``arborist(...)`` call live.

    ``engines``: the contexts of this process (tree ``t`` of this process's shard is
    local tree ``l // G`` of ``engines[l % G]``, ``l`` = position in the shard);
    ``shard``: global indices fitted by this process; ``owner``: rank owning every
    global tree (all zeros without a process group)."""

    def __init__(self, engines, shard, owner, n_trees, batch_bytes=512 << 20, max_batch=256):
        self.engines = list(engines)
        self.shard = np.asarray(shard, dtype=np.int64)
        self.owner = np.asarray(owner, dtype=np.int64)
        self.n_trees = int(n_trees)
        self._pos = {int(g): i for i, g in enumerate(self.shard)}
        n, m, _, _ = self.engines[0].dims()
        K = self.engines[0].n_clones
        self.per_tree_bytes = 8 * (n * K + m * (K - 1))
        self.batch = int(max(1, min(max_batch, batch_bytes // max(1, self.per_tree_bytes))))
        self._dev_batch = {}                   # engine index -> first local tree of the batch cached on its GPU
        self.fetches = 0                       # batches made resident (tests / profiles)

    # -- local trees -----------------------------------------------------------------------------
    def locate(self, idx):
        """(engine index, local tree) of a tree of this process, else ``None``."""
        i = self._pos.get(int(idx))
        if i is None:
            return None
        G = len(self.engines)
        return i % G, i // G

    def _local_count(self, g):
        G = len(self.engines)
        return (len(self.shard) - g + G - 1) // G

    def _make_resident(self, g, l):
        """The batch of local trees around ``l`` on its GPU: a no-op when the posteriors were kept there, one
        batched re-fit otherwise (cached on the device until another batch is asked for).  The host then copies
        single trees out of it, so no host buffer ever holds more than one tree."""
        l0 = (l // self.batch) * self.batch
        if self._dev_batch.get(g) != l0:
            self.engines[g].get_fits(l0, min(self.batch, self._local_count(g) - l0), fetch=False)
            self._dev_batch[g] = l0
            self.fetches += 1

    def _fetch_local(self, g, l):
        self._make_resident(g, l)
        return self.engines[g].get_fit(l)

    def get_map(self, idx):
        """Device-computed MAP vectors of a local tree (``None`` for a remote one)."""
        loc = self.locate(idx)
        if loc is None:
            return None
        self._make_resident(*loc)
        return self.engines[loc[0]].get_map(loc[1])

    def get_fit(self, idx):
        """``(q_z, q_y)`` of global tree ``idx`` as fresh arrays.  A tree fitted by another rank is broadcast
        by its owner: under ``torch.distributed`` that is a COLLECTIVE -- every rank has to ask for the same
        remote tree in the same order (the CLI and :func:`write_fits_pickle` do)."""
        from . import distributed

        loc = self.locate(idx)
        rank, ws = distributed.world()
        if ws == 1:
            return self._fetch_local(*loc)
        if loc is not None:
            qz, qy = self._fetch_local(*loc)
        else:
            n, m, _, _ = self.engines[0].dims()
            K = self.engines[0].n_clones
            qz, qy = np.empty((n, K)), np.empty((m, K - 1))
        return distributed.broadcast_fit(qz, qy, int(self.owner[int(idx)]), self.engines[0].device)


class LazyFits(dict):
    """``{tree_idx: TreeFit}`` for tree indices ``0..n-1`` whose values are built by ``factory(idx)`` on first
    access and then kept, so it behaves like the reference's eager dict (``len``, iteration, ``in``, ``items``,
    ``pd.to_pickle``) without building ``n`` objects per call.  :meth:`to_pickle` writes the same file while
    holding one batch of posteriors at a time."""

    def __init__(self, n, factory, store=None):
        super().__init__()
        self._n, self._factory, self._store = int(n), factory, store

    def __missing__(self, key):
        if isinstance(key, (int, np.integer)) and not isinstance(key, bool) and 0 <= int(key) < self._n:
            fit = self._factory(int(key))
            dict.__setitem__(self, int(key), fit)
            return fit
        raise KeyError(key)

    def __len__(self):
        return self._n

    def __iter__(self):
        return iter(range(self._n))

    def __contains__(self, key):
        return isinstance(key, (int, np.integer)) and not isinstance(key, bool) and 0 <= int(key) < self._n

    def keys(self):
        return range(self._n)

    def values(self):
        return (self[i] for i in range(self._n))

    def items(self):
        return ((i, self[i]) for i in range(self._n))

    def get(self, key, default=None):
        return self[key] if key in self else default

    def __repr__(self):
        return f"LazyFits({self._n} trees, {dict.__len__(self)} materialised)"

    def __eq__(self, other):
        return dict(self.items()) == other

    def __reduce__(self):
        # an ordinary dict in the pickle (pd.to_pickle(all_fits) as in A.py:1187-1188)
        return (dict, (), None, None, iter(self.items()))

    def fresh(self, idx):
        """A TreeFit that is NOT kept in the dict (streaming writers)."""
        return self._factory(int(idx))

    def to_pickle(self, path, protocol=3):
        write_fits_pickle(self, path, protocol=protocol)


class _NullSink:
    def write(self, b):
        return len(b)


def write_fits_pickle(fits, path, protocol=3):
    """Write ``{idx: TreeFit}`` as one pickle, one batch of posteriors in memory at a time.

    The stream is assembled by hand: ``PROTO, EMPTY_DICT, MARK, (key value)*, SETITEMS, STOP``, where every
    ``key`` / ``value`` piece is a ``dump`` of ONE ``pickle.Pickler`` (minus its PROTO and STOP) whose memo persists between pieces, so the index
    dicts shared by all fits are written once and referenced afterwards (as ``pickle.dump`` of the eager dict
    would); after each piece the memo entries that only that piece's arrays needed are dropped, which is what
    keeps the big arrays from staying alive.  Protocol 2 or 3 (explicit memo indices; 4+ numbers the memo
    implicitly, which the trimming would break).  ``path`` may carry a compression suffix like ``pd.to_pickle``.

    Under ``torch.distributed`` every rank must call this (remote posteriors are broadcast); rank 0 writes."""
    import io

    from pandas.io.common import get_handle

    from . import distributed

    if protocol not in (2, 3):
        raise ValueError("write_fits_pickle: protocol 2 or 3")
    rank, _ = distributed.world()
    n = len(fits)
    lazy = isinstance(fits, LazyFits)
    handle = get_handle(path, "wb", compression="infer", is_text=False) if rank == 0 else None
    writer = handle is not None                   # the other ranks only take part in the broadcasts: nothing is pickled
    out = handle.handle if writer else _NullSink()  # or buffered there (the file is 2.7 GB at config 3)
    buf = io.BytesIO()
    pk = pickle.Pickler(buf, protocol=protocol)
    shared_ids = None
    try:
        out.write(pickle.PROTO + bytes([protocol]) + pickle.EMPTY_DICT)
        for start in range(0, n, 1000):                          # SETITEMS batches, like the C pickler
            out.write(pickle.MARK)
            for idx in range(start, min(n, start + 1000)):
                fit = fits.fresh(idx) if lazy else fits[idx]
                fit.q_z  # noqa: B018 - materialise (collective under torch.distributed)
                if not writer:
                    del fit
                    continue
                for obj in (idx, fit):                           # key, then value, left on the unpickler's stack
                    buf.seek(0)
                    buf.truncate()
                    pk.dump(obj)                                 # PROTO <p> ... STOP
                    out.write(buf.getvalue()[2:-1])
                memo = pk.memo.copy()
                if shared_ids is None:
                    # objects every later fit shares with this one (the four index dicts and what hangs off them)
                    shared = [fit.cell_to_idx, fit.snv_to_idx, fit.clone_to_idx, fit.cluster_to_idx,
                              fit.idx_to_clone, fit.idx_to_cluster]
                    shared_ids = {id(o) for o in shared}
                    keep_below = max((v[0] for k, v in memo.items() if k in shared_ids), default=-1) + 1
                    # everything memoised up to the last shared dict stays (class, dict keys / values of those dicts)
                    base_memo = {k: v for k, v in memo.items() if v[0] < keep_below}
                    # later pieces restart numbering right after the kept entries
                    base_memo = dict(sorted(base_memo.items(), key=lambda kv: kv[1][0]))
                    if [v[0] for v in base_memo.values()] != list(range(len(base_memo))):
                        base_memo = {}                            # not a dense prefix: share nothing, stay correct
                pk.memo = dict(base_memo)
                del fit
            out.write(pickle.SETITEMS)
        out.write(pickle.STOP)
    finally:
        if handle is not None:
            handle.close()
