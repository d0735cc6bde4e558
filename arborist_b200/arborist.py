"""Python entry points: ``arborist(...)``, the 13-argument per-tree seam and the CLI.

Drop-in for ``/root/reference/src/arborist/arborist.py`` ("A.py"):

* ``arborist(tree_list, read_counts, snv_clusters, alpha, max_iter, tolerance,
  gamma, add_normal, threads, verbose)`` -- same signature, defaults, return
  triple and ``ValueError`` as A.py:906-1130.  The reference's sequential tree
  loop (A.py:1089-1126) collapses into ONE batched call into the CUDA library.
* ``run_variational_inference(presence, ll_cell_sort, ll_snv_sort, cell_idx,
  snv_idx, n_cells, n_snvs, n_clones, q_y_init, cell_ptr, snv_ptr, max_iter,
  tolerance)`` -- the function-valued seam ``run = ...`` at A.py:1076, same
  argument meaning and return triple as A.py:679-816.
* ``main()`` / ``parse_arguments()`` -- the ``arborist`` console script with the
  reference's flags (A.py:26-142, A.py:1133-1191).

Host code here is orchestration only (id factorisation, tree validation,
result wrapping).  Every numeric step runs in ``csrc/`` on the GPU; there is no
CPU fallback -- a missing library or device raises.
"""
from __future__ import annotations

import argparse
import logging
import os
import weakref

import numpy as np
import pandas as pd

from .treefit import TreeFit
from .utils import read_read_counts, read_trees, visualize_tree

__all__ = ["arborist", "run_variational_inference", "run_simple_max_likelihood", "main",
           "parse_arguments", "add_normal_clone", "tree_to_clone_set", "prepare_trees", "prepare_tree_shard",
           "ingest_read_counts", "clear_seam_cache"]

logging.basicConfig(
    format="{asctime} - {levelname} - {message}",
    style="{",
    datefmt="%Y-%m-%d %H:%M:%S",
    level=logging.INFO,
)


# ---------------------------------------------------------------------------
# tree handling (host): A.py:860-903, A.py:1009-1028, A.py:491-520
# ---------------------------------------------------------------------------
def _digraph_edge_order(edge_list, extra=None):
    """Edges in the order ``list(nx.DiGraph(edge_list).edges)`` yields them
    (grouped by source in node first-appearance order, duplicates collapsed);
    the reference stores that order in ``TreeFit.tree`` after ``--add-normal``."""
    nodes, succ = {}, {}
    for u, v in edge_list:
        nodes.setdefault(u, None)
        nodes.setdefault(v, None)
        succ.setdefault(u, {}).setdefault(v, None)
    if extra is not None:
        u, v = extra
        nodes.setdefault(u, None)
        nodes.setdefault(v, None)
        succ.setdefault(u, {}).setdefault(v, None)
    return [(u, v) for u in nodes for v in succ.get(u, ())], list(nodes)


def _add_normal_one(edges):
    """One tree of :func:`add_normal_clone`: node / successor order as ``nx.DiGraph(edges)`` keeps it, then the
    extra edge ``normal -> root`` last (``list(tree.edges)`` after ``tree.add_edge``, A.py:876-880)."""
    nodes, succ, kids = {}, {}, set()
    for u, v in edges:
        nodes.setdefault(u, None)
        nodes.setdefault(v, None)
        succ.setdefault(u, {}).setdefault(v, None)
        kids.add(v)
    root = next(n for n in nodes if n not in kids)
    normal = min(nodes) - 1
    nodes.setdefault(normal, None)
    succ.setdefault(normal, {}).setdefault(root, None)
    return [(u, v) for u in nodes for v in succ.get(u, ())]


def add_normal_clone(candidate_trees):
    """Prepend a normal clone ``min(nodes) - 1`` above each tree's root (A.py:860-882)."""
    return [_add_normal_one(edges) for edges in candidate_trees]


def tree_to_clone_set(tree):
    """Set of clone labels in an edge list (A.py:885-903)."""
    return {x for e in tree for x in e}


def _edge_array(trees):
    """``[T, E, 2]`` int64 array of a regular candidate set (every tree the same number of integer edges), else None."""
    try:
        arr = np.asarray(trees, dtype=np.int64)
    except (ValueError, TypeError, OverflowError):
        return None
    if arr.ndim != 3 or arr.shape[2] != 2 or arr.shape[1] == 0:
        return None
    return arr


def _parents_from_array(arr, clones):
    """Parent vectors ``[T, K]`` (-1 for roots) of the trees in ``arr`` over the sorted label list ``clones``.
    Returns ``(parents, is_forest)``; raises the reference's ValueError when a tree's label set differs."""
    T, K = arr.shape[0], len(clones)
    carr = np.asarray(clones, dtype=np.int64)
    idx = np.searchsorted(carr, arr).clip(0, K - 1)
    if not np.array_equal(carr[idx], arr):
        raise ValueError("All trees must have the same set of clones.")
    seen = np.zeros((T, K), dtype=bool)
    rows = np.arange(T)[:, None]
    seen[rows, idx[:, :, 0]] = True
    seen[rows, idx[:, :, 1]] = True
    if not seen.all():
        raise ValueError("All trees must have the same set of clones.")
    parents = np.full((T, K), -1, dtype=np.int32)
    pu, cv = idx[:, :, 0].astype(np.int32), idx[:, :, 1]
    parents[rows, cv] = pu
    is_forest = True
    # a child listed under two different parents is not a forest
    n_par = np.zeros((T, K), dtype=np.int32)
    np.add.at(n_par, (np.broadcast_to(rows, cv.shape), cv), 1)
    if (n_par > 1).any():
        # tolerate exact duplicate edges, reject genuine multi-parent nodes
        is_forest = bool((parents[rows, cv] == pu).all())
    return parents, is_forest


def _acyclic(parents):
    """After K pointer jumps every node of a forest has reached a root."""
    K = parents.shape[1]
    cur = parents.copy()
    for _ in range(K):
        cur = np.where(cur >= 0, np.take_along_axis(parents, cur.clip(0), axis=1), -1)
    return bool((cur < 0).all())


def _presence_on_host(tree_list, clone_to_idx, cluster_to_idx):
    """General digraphs: descendant closure on the host, like nx.descendants (A.py:516)."""
    T, K = len(tree_list), len(clone_to_idx)
    presence = np.zeros((T, K - 1, K), dtype=np.uint8)
    for t, tree in enumerate(tree_list):
        children = {}
        for u, v in tree:
            children.setdefault(u, []).append(v)
        for cl, ip in cluster_to_idx.items():
            stack, seen_n = [cl], set()
            while stack:
                x = stack.pop()
                if x in seen_n:
                    continue
                seen_n.add(x)
                presence[t, ip, clone_to_idx[x]] = 1
                stack.extend(children.get(x, ()))
    return presence


def _index_trees(tree_list, clones, clone_to_idx, cluster_to_idx):
    """Parent vectors (forests) or genotype matrices (anything else) of fully materialised edge lists."""
    K, T = len(clones), len(tree_list)
    clone_set = set(clones)
    arr = _edge_array(tree_list)
    if arr is not None:
        parents, is_forest = _parents_from_array(arr, clones)
    else:
        parents = np.full((T, K), -1, dtype=np.int32)
        is_forest = True
        for t, tree in enumerate(tree_list):
            if tree_to_clone_set(tree) != clone_set:
                raise ValueError("All trees must have the same set of clones.")
            for u, v in tree:
                iu, iv = clone_to_idx[u], clone_to_idx[v]
                if parents[t, iv] not in (-1, iu):
                    is_forest = False
                parents[t, iv] = iu
    if is_forest and _acyclic(parents):
        return parents, None
    return None, _presence_on_host(tree_list, clone_to_idx, cluster_to_idx)


def _label_maps(tree0):
    """Clone / cluster index maps and the normal clone from the first tree (A.py:1009-1028)."""
    _, nodes0 = _digraph_edge_order(tree0)
    kids0 = {v for _, v in tree0}
    normal = [n for n in nodes0 if n not in kids0][0]
    clones = sorted(tree_to_clone_set(tree0))
    clusters = [c for c in clones if c != normal]
    return (clones, normal, {c: i for i, c in enumerate(clones)}, {c: i for i, c in enumerate(clusters)})


def prepare_trees(tree_list):
    """Validate the candidate set and index it (A.py:1009-1028).

    Returns ``(clone_to_idx, cluster_to_idx, parents, presence, normal_idx)``; exactly one of
    ``parents`` (int32 ``[T, K]``, -1 for roots; device builds the closure) and
    ``presence`` (uint8 ``[T, K-1, K]``; for inputs that are not forests) is set.
    """
    clones, normal, clone_to_idx, cluster_to_idx = _label_maps(tree_list[0])
    parents, presence = _index_trees(tree_list, clones, clone_to_idx, cluster_to_idx)
    return clone_to_idx, cluster_to_idx, parents, presence, clone_to_idx[normal]


def prepare_tree_shard(tree_list, indices, add_normal=False):
    """:func:`prepare_trees` for the trees ``indices`` of the candidate set only (the shard of this process), with
    ``--add-normal`` (A.py:860-882, 1004-1007) folded in: for a regular integer candidate set the normal clone is
    added arithmetically to the parent vectors and no per-tree edge list is built; host work is proportional to
    the shard, not to the whole candidate set.  Label maps always come from tree 0, as in the reference."""
    tree0 = _add_normal_one(tree_list[0]) if add_normal else tree_list[0]
    clones, normal, clone_to_idx, cluster_to_idx = _label_maps(tree0)
    normal_idx = clone_to_idx[normal]
    local = [tree_list[int(i)] for i in indices]
    if not local:
        return clone_to_idx, cluster_to_idx, np.zeros((0, len(clones)), np.int32), None, normal_idx
    if add_normal:
        arr = _edge_array(local)
        raw0 = tree_to_clone_set(tree_list[0])
        if arr is not None and all(isinstance(c, (int, np.integer)) for c in raw0) and min(raw0) - 1 == normal:
            raw = sorted(raw0)                                   # = clones without the normal, which sorts first
            par_raw, is_forest = _parents_from_array(arr, raw)
            one_root = bool(((par_raw < 0).sum(axis=1) == 1).all())
            if is_forest and one_root and _acyclic(par_raw):
                parents = np.empty((len(local), len(clones)), dtype=np.int32)
                parents[:, 0] = -1                               # the normal clone
                parents[:, 1:] = np.where(par_raw >= 0, par_raw + 1, 0)
                return clone_to_idx, cluster_to_idx, parents, None, normal_idx
        local = add_normal_clone(local)
    parents, presence = _index_trees(local, clones, clone_to_idx, cluster_to_idx)
    return clone_to_idx, cluster_to_idx, parents, presence, normal_idx


def _prior_labels(snv_clusters, snv_ids, cluster_to_idx):
    """Input-cluster index per SNV row, -1 => uniform prior row (A.py:547-567).

    Kept quirks: the LAST row wins for a repeated SNV (``dict(zip(...))``); a label
    that is falsy (``0``, ``None``, ``""``) or not a tree node gives the uniform
    row; SNVs absent from the clustering are reported on stdout (A.py:558).
    """
    cl = snv_clusters.drop_duplicates("snv", keep="last")
    lab = pd.Series(cl["cluster"].to_numpy(), index=cl["snv"].to_numpy())
    ids = pd.Index(snv_ids)
    found = ids.isin(lab.index)
    for s in ids[~found]:
        print(f"SNV {s} not initialized with cluster assignment.")
    got = lab.reindex(ids)
    vals = got.to_numpy()
    out = np.full(len(ids), -1, dtype=np.int32)
    mapped = got.map(cluster_to_idx).to_numpy(dtype=np.float64, na_value=np.nan)
    ok = found & ~np.isnan(mapped)
    if vals.dtype.kind in "iufb":
        ok &= vals != 0
    else:
        ok &= np.fromiter((bool(v) if v is not None else False for v in vals), dtype=bool,
                          count=len(vals))
    out[ok] = mapped[ok].astype(np.int32)
    return out


def ingest_read_counts(eng, read_counts, snv_clusters, cluster_to_idx, alpha, gamma):
    """A.py:1034-1069 on the device: upload the long-format table as integer codes, build the
    sparse matrix and the prior there.  Returns ``(cell_ids, snv_ids)`` in row/column order.

    The first-appearance factorisation of the two id columns (A.py:851-855, on the UNFILTERED
    table) runs on the device when the ids are int64, on the host (pandas) otherwise; the zero-alt filter (A.py:1034-1041), re-indexing, dictionary
    coding and the cell-major / SNV-major layouts are built by ``arb_ingest_coo``.
    """
    if isinstance(read_counts, (str, os.PathLike)):
        # a file instead of a DataFrame: integer CSVs are parsed on the device (arb_csv_*), anything else (text ids,
        # quoted fields, Parquet / Feather / npz) goes through the host reader first
        from .engine import ArboristError
        path = os.fspath(read_counts)
        read_counts = None
        if path.lower().endswith(".csv"):
            try:
                cell_uniques, snv_uniques, _ = eng.ingest_csv(path, alpha)
            except ArboristError as e:
                if e.status != -5:                      # only ARB_ERR_DECLINED (not a plain integer table) falls back
                    raise
                read_counts = read_read_counts(path)
        else:
            read_counts = read_read_counts(path)
    cols = [] if read_counts is None else [read_counts[c].to_numpy() for c in ("cell", "snv", "alt", "total")]
    if read_counts is None:
        pass
    elif all(a.dtype == np.int64 for a in cols):
        # integer ids: the factorisation runs on the device too (arb_ingest_ids)
        cell_uniques, snv_uniques = eng.ingest_ids(*cols, alpha)
    else:
        snv_codes, snv_uniques = pd.factorize(cols[1], use_na_sentinel=True)
        cell_codes, cell_uniques = pd.factorize(cols[0], use_na_sentinel=False)
        eng.ingest_coo(cell_codes, snv_codes, cols[2], cols[3], len(cell_uniques), len(snv_uniques), alpha)
    _install_scipy_table(eng, alpha)
    cell_order, snv_order = eng.orders()
    cell_ids = np.asarray(cell_uniques)[cell_order]
    snv_ids = np.asarray(snv_uniques)[snv_order]
    eng.set_prior_labels(_prior_labels(snv_clusters, snv_ids, cluster_to_idx),
                         len(cluster_to_idx), gamma)
    return cell_ids, snv_ids


# ---------------------------------------------------------------------------
# the public entry point
# ---------------------------------------------------------------------------
def _pick_devices(device, devices, n_trees, work=0.0):
    """GPUs one ``arborist(...)`` call drives.  Under a ``torch.distributed`` group: the one GPU of this rank.
    Otherwise ``devices`` ("all", a count, or a list of ordinals) or ``device``; the default is every visible GPU
    once the job is large enough to pay for them -- every extra GPU costs a CUDA context (~0.2 s) and an ingest of
    the read-count table, so a fit that one GPU finishes in about a second (config 3: table rows x clones x trees
    = 4.8e10, 0.66 s) stays on one; ``work`` is that product."""
    from . import distributed
    from .engine import device_count

    _, ws = distributed.world()
    if ws > 1 or device is not None:
        return [device]                         # None -> LOCAL_RANK inside Engine
    n = device_count()
    if devices is None:
        env = os.environ.get("ARBORIST_DEVICES", "")
        devices = env if env else ("all" if n_trees >= 32 * max(1, n) and work >= 1.5e11 else 1)
    if isinstance(devices, str):
        # "all", a bare integer = a COUNT (like devices=2), or a comma-separated list of ordinals ("3," = GPU 3 alone)
        txt = devices.strip()
        if txt == "all":
            devices = n
        elif "," in txt:
            devices = [int(x) for x in txt.split(",") if x.strip() != ""]
        else:
            devices = int(txt)
    if isinstance(devices, (int, np.integer)):
        devices = list(range(max(1, min(int(devices), n))))
    devices = [int(d) for d in devices]
    if not devices:
        raise ValueError("arborist: empty device list")
    if n > 0:                                   # without any GPU the engine reports "no usable CUDA device" itself
        for d in devices:
            if not 0 <= d < n:
                raise ValueError(f"arborist: CUDA device {d} requested but this process sees {n} (ordinals 0..{n - 1})")
    return devices[: max(1, n_trees)]


def arborist(
    tree_list: list,
    read_counts: pd.DataFrame,
    snv_clusters: pd.DataFrame,
    alpha: float = 0.001,
    max_iter: int = 10,
    tolerance: float = 1,
    gamma=0.7,
    add_normal=False,
    threads=10,
    verbose: bool = False,
    *,
    device=None,
    devices=None,
    keep_posteriors: str = "auto",
) -> tuple:
    """Rank candidate SNV clone trees with low-pass scDNA-seq read counts.

    Same contract as the reference (A.py:906-1130): returns ``(likelihoods,
    best_fit, all_tree_fits)`` -- ``{tree_idx: ELBO}``, the ``TreeFit`` of the first
    tree attaining the maximum ELBO, and ``{tree_idx: TreeFit}`` (a ``dict`` whose
    values are built when first looked at, :class:`arborist_b200.results.LazyFits`).

    ``threads`` is accepted for compatibility and ignored (the reference passes it
    to ``numba.set_num_threads``, A.py:1002; nothing here runs on host threads).
    Keyword-only extras: ``device`` (one CUDA ordinal), ``devices`` ("all", a count -- also as
    a bare integer string -- or a list of ordinals, as a string "0,2" or "3,": the candidate set is spread over those GPUs by THIS process,
    ``arb_fit_trees_multi``; default: every visible GPU once the job is worth their
    start-up cost -- at least 32 trees per GPU and table rows x clones x trees >= 1.5e11, i.e.
    about 2 s of one GPU; environment override ``ARBORIST_DEVICES``) and ``keep_posteriors``
    ("auto" | "all" | "none": whether every tree's best ``q_z``/``q_y`` stay resident
    on the GPU or are re-fitted in batches on demand).
    Under an initialised ``torch.distributed`` group (one process per GPU) the candidate
    set is sharded over ranks instead and the ELBOs gathered, see
    :mod:`arborist_b200.distributed`.
    """
    import gc
    import time
    from concurrent.futures import ThreadPoolExecutor

    from . import distributed
    from .engine import Engine, fit_multi
    from .results import FitStore, LazyFits

    logger = logging.getLogger() if verbose else None

    def log(msg):
        if logger is not None:
            logger.info(msg)

    marks = [("start", time.perf_counter())]         # host wall-clock of the stages, kept in arborist.last_stats

    def mark(name):
        marks.append((name, time.perf_counter()))

    T = len(tree_list)
    shard = distributed.shard_indices(T)
    if add_normal:
        log("Appending normal clone to candidate trees...")
    # validation + parent vectors of this process's shard only; --add-normal folded in (no per-tree edge lists)
    err = None
    try:
        clone_to_idx, cluster_to_idx, parents, presence, normal_idx = prepare_tree_shard(tree_list, shard, add_normal)
    except Exception as e:                              # anything: the other ranks are about to enter a collective
        err = e
    if not distributed.all_agree(err is None, device):  # every rank raises, or none
        raise err if err is not None else ValueError("All trees must have the same set of clones.")
    K = len(clone_to_idx)
    mark("trees")

    log("Removing SNVs with 0 variant reads across all cells...")
    if isinstance(read_counts, (str, os.PathLike)):
        rows = os.path.getsize(read_counts) / 15.0           # ~15 bytes per CSV row
    else:
        rows = float(len(read_counts))
    devs = _pick_devices(device, devices, len(shard), rows * K * len(shard))
    engines = [Engine(device=d) for d in devs]
    mark("context")
    log("Caching log-likelihoods for presence/absence...")
    log("Initializing the SNV cluster assignment prior...")
    if len(engines) == 1:
        cell_ids, snv_ids = ingest_read_counts(engines[0], read_counts, snv_clusters, cluster_to_idx, alpha, gamma)
    else:
        # one ingest per GPU, side by side (ctypes calls release the interpreter lock)
        with ThreadPoolExecutor(len(engines)) as ex:
            got = list(ex.map(lambda e: ingest_read_counts(e, read_counts, snv_clusters, cluster_to_idx, alpha, gamma),
                              engines))
        cell_ids, snv_ids = got[0]
    mark("ingest")
    cell_to_idx = dict(zip(cell_ids.tolist(), range(len(cell_ids))))
    snv_to_idx = dict(zip(snv_ids.tolist(), range(len(snv_ids))))
    mark("index_maps")

    if verbose:
        log("Starting Arborist...")
        log(f"Number of candidate trees: {len(tree_list[-1]) + (1 if add_normal else 0)}")   # sic: A.py:1080 logs an edge count
        log(f"Number of clones: {K}")
        log(f"Number of SNV clusters: {K - 1}")
        log(f"Number of cells: {len(cell_ids)}")
        log(f"Number of SNVs: {len(snv_ids)}")

    if len(engines) == 1 or parents is None:
        engines = engines[:1]
        elbo, iters = engines[0].fit(parents=parents, presence=presence, n_clones=K, normal_idx=normal_idx,
                                     max_iter=max_iter, tolerance=tolerance, keep_posteriors=keep_posteriors)
    else:
        elbo, iters = fit_multi(engines, parents, K, normal_idx=normal_idx, max_iter=max_iter,
                                tolerance=tolerance, keep_posteriors=keep_posteriors)
    mark("fit")
    elbo_all, owner = distributed.gather_elbos(elbo, shard, T, device=engines[0].device)
    mark("gather")

    store = FitStore(engines, shard, owner, T)
    if add_normal:
        def tree_of(idx):
            return _add_normal_one(tree_list[idx])
    else:
        def tree_of(idx):
            return tree_list[idx]
    elbo_list = elbo_all.tolist()

    def make_fit(idx):
        return TreeFit.lazy(None, idx, elbo_list[idx], store.get_fit, cell_to_idx, snv_to_idx, clone_to_idx,
                            cluster_to_idx, map_loader=store.get_map, tree_loader=tree_of)

    all_tree_fits = LazyFits(T, make_fit, store)
    likelihoods = dict(enumerate(elbo_list))
    if verbose:
        for idx, e in enumerate(elbo_list):
            log(f"Tree {idx} fit wtih ELBO: {e:.2f}")
    log(f"Done fitting all {T} candidate trees!")
    # strict >: the first maximum wins, NaN never does (A.py:1124)
    cand = np.where(np.isnan(elbo_all), -np.inf, elbo_all)
    best_idx = int(np.argmax(cand)) if T else -1
    if T == 0 or not cand[best_idx] > -np.inf:
        # the reference dies with UnboundLocalError here (A.py:1124-1130); say why
        raise UnboundLocalError("no candidate tree has a finite ELBO (all NaN/-inf); "
                                "cannot select best_fit")
    best_fit = all_tree_fits[best_idx]
    mark("treefits")
    stages = {b[0]: round(1e3 * (b[1] - a[1]), 2) for a, b in zip(marks, marks[1:])}
    arborist.last_stats = dict(iters=iters, engine=weakref.ref(engines[0]), engines=[weakref.ref(e) for e in engines],
                               shard=shard, stages_ms=stages, store=weakref.ref(store), devices=[e.device for e in engines])
    # Opt-in (ARBORIST_GC_FREEZE=1; bench.py does the same itself): the start-up heap of the interpreter (torch / pandas
    # module objects, never garbage) makes every few calls pay a ~50 ms generation-2 collection over objects that
    # cannot die.  A long-running service may take it out of the collector once; collect first so that no pending
    # cyclic garbage of the host application is frozen with it.  Off by default: the collector belongs to the caller.
    if not arborist._frozen and os.environ.get("ARBORIST_GC_FREEZE", "0") == "1":
        arborist._frozen = True
        gc.collect()
        gc.freeze()
    return likelihoods, best_fit, all_tree_fits


arborist._frozen = False
arborist.last_stats = None


def _install_scipy_table(eng, alpha):
    """Replace the device-computed log-likelihood table with scipy's values.

    The reference evaluates ``scipy.stats.binom.logpmf`` per row (A.py:838-846);
    the device dictionary-codes the distinct ``(alt, total)`` pairs, so evaluating
    scipy on that small table reproduces the reference's doubles bit for bit.
    Without scipy the device's own lgamma-based table (<= 1e-14 relative) stays.
    """
    try:
        from scipy.stats import binom
    except ImportError:      # pragma: no cover - scipy is a dependency of the reference too
        return
    alt, total = eng.count_pairs()
    lp = binom.logpmf(alt, total, 0.5 - alpha + 0.5 * alpha / 3)
    la = binom.logpmf(alt, total, alpha / 3)
    eng.set_loglik_table(lp, la)


# ---------------------------------------------------------------------------
# the per-tree seam (A.py:1076 ``run = run_variational_inference``)
# ---------------------------------------------------------------------------
_SEAM_CACHE = {}


def _fingerprint(a):
    """Content check of a cached input over ALL of its bytes, so that an array refilled in place between calls -- a
    gamma sweep reusing one q_y_init buffer, an edit or a swap of a few log-likelihoods -- is noticed wherever it
    falls: the 64-bit words are laid out as rows of 4096 and both the row sums and the column sums (wrapping) are
    kept, plus the bytes that do not fill a row verbatim.  Changing one word changes a row sum; exchanging two words
    changes a row sum or a column sum.  Two memory-speed passes: about 15 ms for the 64 MB of log-likelihoods at
    config 3 (zlib.crc32 takes 35 ms)."""
    raw = np.ascontiguousarray(a).reshape(-1).view(np.uint8)
    rows = raw.shape[0] // (8 * 4096)
    body = raw[: rows * 8 * 4096].view(np.uint64).reshape(rows, 4096)
    sums = (np.add.reduce(body, axis=1, dtype=np.uint64).tobytes(), np.add.reduce(body, axis=0, dtype=np.uint64).tobytes())
    return (a.__array_interface__["data"][0], a.shape, a.dtype.str, sums, raw[rows * 8 * 4096:].tobytes())


def clear_seam_cache():
    """Drop the device matrix the per-tree seam keeps between calls (and the arrays it pins)."""
    hit = _SEAM_CACHE.pop("e", None)
    if hit is not None:
        hit["engine"].close()


def _seam_engine(orig, ll_cs, snv_idx, cell_ptr, n_cells, n_snvs, q_y_init):
    """The engine behind ``run_variational_inference``: the reference's tree loop calls the seam once per tree with
    the same arrays, so the device matrix is kept between calls.  The key is taken from the CALLER's arrays
    (``orig``: before any dtype / contiguity conversion, so int64 pointers do not defeat the cache) plus a content
    fingerprint; only the current prior is kept."""
    from .engine import Engine

    key = tuple(_fingerprint(np.asarray(a)) for a in orig[:3]) + (n_cells, n_snvs)
    hit = _SEAM_CACHE.get("e")
    if hit is None or hit["key"] != key:
        clear_seam_cache()
        eng = Engine()
        eng.ingest_csr(cell_ptr, snv_idx, ll_cs, n_cells, n_snvs)
        hit = dict(key=key, engine=eng, pinned=orig[:3], prior=None)     # pins the arrays: addresses cannot be recycled
        _SEAM_CACHE["e"] = hit
    pk = _fingerprint(np.asarray(orig[3]))
    if hit["prior"] is None or hit["prior"][0] != pk:
        hit["engine"].set_prior(q_y_init)
        hit["prior"] = (pk, orig[3])
    return hit["engine"]


def _run_seam(mode, presence, ll_cs, ll_ss, cell_idx, snv_idx, n_cells, n_snvs, n_clones,
              q_y_init, cell_ptr, snv_ptr, max_iter, tolerance):
    orig = (ll_cs, snv_idx, cell_ptr, q_y_init)
    ll_cs = np.ascontiguousarray(ll_cs, dtype=np.float64)
    snv_idx = np.ascontiguousarray(snv_idx, dtype=np.int32)
    cell_ptr = np.ascontiguousarray(cell_ptr, dtype=np.int32)
    q_y_init = np.ascontiguousarray(q_y_init, dtype=np.float64)
    eng = _seam_engine(orig, ll_cs, snv_idx, cell_ptr, int(n_cells), int(n_snvs), q_y_init)
    pres = (np.asarray(presence) != 0).astype(np.uint8)[None]
    elbo, _ = eng.fit(presence=pres, n_clones=int(n_clones), max_iter=int(max_iter),
                      tolerance=float(tolerance), keep_posteriors="all", mode=mode)
    q_z, q_y = eng.get_fit(0)
    return float(elbo[0]), q_z, q_y


def run_variational_inference(presence, log_like_matrix_cell_sort, log_like_matrix_snv_sort,
                              cell_idx, snv_idx, n_cells, n_snvs, n_clones, q_y_init, cell_ptr,
                              snv_ptr, max_iter=10, tolerance=1):
    """CAVI fit of one tree; arguments and return triple as A.py:679-816.

    The CSC copy (``log_like_matrix_snv_sort``, ``cell_idx``, ``snv_ptr``) is accepted
    and ignored: the device derives its own SNV-major layout from the CSR arrays.
    The device matrix is cached across calls on the same arrays (the reference's
    tree loop calls this once per tree with identical data)."""
    return _run_seam("vi", presence, log_like_matrix_cell_sort, log_like_matrix_snv_sort,
                     cell_idx, snv_idx, n_cells, n_snvs, n_clones, q_y_init, cell_ptr, snv_ptr,
                     max_iter, tolerance)


def run_simple_max_likelihood(presence, log_like_matrix_cell_sort, log_like_matrix_snv_sort,
                              cell_idx, snv_idx, n_cells, n_snvs, n_clones, q_y_init, cell_ptr,
                              snv_ptr, max_iter=10, tolerance=1):
    """Hard-assignment scorer with the seam's signature (A.py:598-676)."""
    return _run_seam("ml", presence, log_like_matrix_cell_sort, log_like_matrix_snv_sort,
                     cell_idx, snv_idx, n_cells, n_snvs, n_clones, q_y_init, cell_ptr, snv_ptr,
                     max_iter, tolerance)


# ---------------------------------------------------------------------------
# CLI (A.py:26-142, A.py:1133-1191)
# ---------------------------------------------------------------------------
def parse_arguments(argv=None) -> argparse.Namespace:
    p = argparse.ArgumentParser(
        description="Arborist: a method to rank SNV clonal trees using scDNA-seq data.")
    p.add_argument("-R", "--read_counts", required=True,
                   help="Path to read counts CSV file with columns 'snv', 'cell',  'total', 'alt'")
    p.add_argument("-Y", "--snv-clusters", required=True,
                   help="Path to SNV clusters CSV file with unlabeled columns 'snv', 'cluster'. "
                        "The order of columns matters")
    p.add_argument("-T", "--trees", required=True,
                   help="Path to file containing all candidate trees to be ranked.")
    p.add_argument("--edge-delim", type=str, default=" ",
                   help="edge delimiter in candidate tree file.")
    p.add_argument("--add-normal", action="store_true",
                   help="flag to add a normal clone if input trees do not already contain them")
    p.add_argument("--alpha", type=float, default=0.001, help="Per base sequencing error")
    p.add_argument("--max-iter", type=int, default=25, help="max number of iterations")
    p.add_argument("--prior", type=float, default=0.7,
                   help="prior (gamma) on input SNV cluster assignment")
    p.add_argument("--pickle", type=str,
                   help="path to where all pickled tree fits should be saved.")
    p.add_argument("-d", "--draw", type=str,
                   help="Path to where the tree visualization should be saved")
    p.add_argument("-t", "--tree", help="Path to save the top ranked tree as a txt file.")
    p.add_argument("--ranking", type=str,
                   help="Path to where tree ranking output should be saved")
    p.add_argument("--cell-assign", type=str,
                   help="Path to where the MAP cell-to-clone labels should be saved")
    p.add_argument("--snv-assign", type=str,
                   help="Path to where the MAP SNV-to-cluster labels should be saved.")
    p.add_argument("--q_y", type=str,
                   help="Path to where the approximate SNV posterior should be saved")
    p.add_argument("--q_z", type=str,
                   help="Path to where the approximate cell posterior should be saved")
    p.add_argument("-j", "--threads", default=10, type=int, help="Number of threads to use")
    p.add_argument("-v", "--verbose", help="Print verbose output", action="store_true")
    return p.parse_args(argv)


def main(argv=None):
    args = parse_arguments(argv)
    read_counts = args.read_counts     # the file itself: CSV as upstream (parsed on the device when its columns are
                                       # integers, by pandas otherwise); also Parquet / Feather / npz
    snv_clusters = pd.read_csv(args.snv_clusters, header=None, names=["snv", "cluster"])
    candidate_trees = read_trees(args.trees, sep=args.edge_delim)

    elbos, tfit, all_fits = arborist(
        candidate_trees, read_counts, snv_clusters, alpha=args.alpha, verbose=args.verbose,
        max_iter=args.max_iter, gamma=args.prior, add_normal=args.add_normal,
        threads=args.threads)

    if args.draw:
        visualize_tree(tfit.tree, output_file=args.draw)

    elbo_df = pd.DataFrame({"tree_idx": list(elbos.keys()), "elbo": list(elbos.values())})
    elbo_df = elbo_df.sort_values(by="elbo", ascending=False)
    if args.ranking:
        elbo_df.to_csv(args.ranking, index=False)
    if args.cell_assign:
        tfit.map_assign_z().to_csv(args.cell_assign, index=False)
    if args.snv_assign:
        tfit.map_assign_y().to_csv(args.snv_assign, index=False)
    if args.q_z:
        tfit.q_z_df().to_csv(args.q_z, index=False)
    if args.q_y:
        tfit.q_y_df().to_csv(args.q_y, index=False)
    if args.tree:
        tfit.save_tree(args.tree)
    if args.pickle:
        # the same file pd.to_pickle(all_fits, ...) writes (A.py:1187-1188), one batch of posteriors at a time
        all_fits.to_pickle(args.pickle)

    print("\n---------------------Arborist complete!---------------------\n")
    print(tfit)
