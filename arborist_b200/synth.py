"""Synthetic read-count / clustering / candidate-tree generator.

The reference's own tests draw their inputs from an external simulator that is
not vendored (``/root/reference/tests/conftest.py:1-53``), so this module
carries a small generator that follows the reference's likelihood model
(``src/arborist/arborist.py:838-846``): a cell observes an SNV with a fixed
probability, ``total = 1 + Poisson``, ``alt ~ Binomial(total, p)`` with
``p = 0.5 - a + 0.5 a / 3`` when the SNV's cluster is an ancestor-or-self of the
cell's clone and ``p = a / 3`` otherwise.

Everything is numpy ``default_rng(seed)`` (PCG64) so the same seed gives the
same bytes on every machine.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

__all__ = ["SynthConfig", "SynthData", "CONFIGS", "generate", "random_tree", "parents_of"]


@dataclass(frozen=True)
class SynthConfig:
    """One named workload. ``n_clones`` counts the normal clone (K)."""

    name: str
    n_cells: int
    n_snvs: int
    n_clones: int
    n_trees: int
    seed: int
    coverage: float = 0.02
    extra_depth: float = 0.1          # total = 1 + Poisson(extra_depth)
    alpha: float = 0.001
    cluster_error_prob: float = 0.05
    offtree_frac: float = 0.0         # SNVs given a label no tree contains
    deep: bool = False                # chain-like true/candidate trees
    min_proportion: float = 0.05


# BASELINE.json configs 2-5 (config 1 is the bundled example, see tests/).
CONFIGS = {
    "tiny": SynthConfig("tiny", 60, 150, 5, 6, seed=1000, coverage=0.15),
    "small": SynthConfig("small", 300, 900, 6, 12, seed=1010, coverage=0.05),
    "config2": SynthConfig("config2", 1000, 5000, 8, 100, seed=1002),
    "config3": SynthConfig("config3", 10000, 20000, 12, 1000, seed=1003),
    "config4": SynthConfig("config4", 100000, 50000, 20, 10000, seed=1004),
    "config5": SynthConfig("config5", 5000, 10000, 30, 100, seed=1005,
                           offtree_frac=0.30, deep=True),
}


@dataclass
class SynthData:
    config: SynthConfig
    # long-format read counts, sorted by (snv, cell); ids are plain ints
    snv: np.ndarray
    cell: np.ndarray
    alt: np.ndarray
    total: np.ndarray
    # input clustering (snv id, cluster label), one row per SNV
    clust_snv: np.ndarray
    clust_label: np.ndarray
    # candidate trees WITHOUT the normal clone: list of [(parent, child), ...]
    trees: list = field(default_factory=list)
    true_tree_index: int = 0
    cell_clone: np.ndarray | None = None     # generating truth (clone label; 0 = normal)
    snv_cluster: np.ndarray | None = None    # generating truth (cluster label)

    def read_counts_frame(self):
        import pandas as pd

        return pd.DataFrame({"snv": self.snv, "cell": self.cell,
                             "alt": self.alt, "total": self.total})

    def clusters_frame(self):
        import pandas as pd

        return pd.DataFrame({"snv": self.clust_snv, "cluster": self.clust_label})


def random_tree(rng: np.random.Generator, n_clusters: int, deep: bool = False) -> tuple:
    """Random recursive tree on labels 1..n_clusters rooted at 1.

    Returns the parent vector as a tuple ``par[c-2]`` = parent of node ``c`` for
    ``c = 2..n_clusters`` (hashable, used for de-duplication).
    """
    par = []
    for c in range(2, n_clusters + 1):
        if deep and rng.random() < 0.8:
            par.append(c - 1)
        else:
            par.append(int(rng.integers(1, c)))
    return tuple(par)


def _edges(par: tuple) -> list:
    return [(int(p), c) for c, p in enumerate(par, start=2)]


def parents_of(edges, labels):
    """Parent label of each label in ``labels`` (``None`` for the root)."""
    pmap = {c: p for p, c in edges}
    return [pmap.get(l) for l in labels]


def _ancestor_closure(par: tuple, n_clusters: int) -> np.ndarray:
    """anc[p, r] = 1 iff cluster p is an ancestor-or-self of clone r.

    Index 0 is the normal clone (carries no cluster); clusters are 1..n_clusters.
    """
    anc = np.zeros((n_clusters + 1, n_clusters + 1), dtype=bool)
    parent = {1: 0}
    for c, p in enumerate(par, start=2):
        parent[c] = p
    for r in range(1, n_clusters + 1):
        x = r
        while x != 0:
            anc[x, r] = True
            x = parent[x]
    return anc


def generate(cfg: SynthConfig) -> SynthData:
    rng = np.random.default_rng(cfg.seed)
    kc = cfg.n_clones - 1                      # number of clusters
    n, m = cfg.n_cells, cfg.n_snvs

    # --- trees -----------------------------------------------------------
    true_par = random_tree(rng, kc, cfg.deep)
    seen = {true_par}
    others = []
    guard = 0
    while len(others) < cfg.n_trees - 1:
        t = random_tree(rng, kc, cfg.deep)
        guard += 1
        if t in seen and guard < 200 * cfg.n_trees:
            continue
        seen.add(t)
        others.append(t)
    true_idx = int(rng.integers(0, cfg.n_trees))
    pars = others[:true_idx] + [true_par] + others[true_idx:]
    trees = [_edges(p) for p in pars]

    # --- latent assignments ---------------------------------------------
    def floored_dirichlet(k):
        w = rng.dirichlet(np.ones(k))
        w = np.maximum(w, cfg.min_proportion)
        return w / w.sum()

    cell_clone = rng.choice(cfg.n_clones, size=n, p=floored_dirichlet(cfg.n_clones))
    snv_cluster = 1 + rng.choice(kc, size=m, p=floored_dirichlet(kc))

    noisy = snv_cluster.copy()
    flip = rng.random(m) < cfg.cluster_error_prob
    noisy[flip] = 1 + rng.integers(0, kc, size=int(flip.sum()))
    if cfg.offtree_frac > 0:
        off = rng.random(m) < cfg.offtree_frac
        noisy[off] = 9999

    # --- observations ----------------------------------------------------
    anc = _ancestor_closure(true_par, kc)
    p_pres = 0.5 - cfg.alpha + 0.5 * cfg.alpha / 3
    p_abs = cfg.alpha / 3

    # per SNV: Binomial(n, coverage) observed cells, drawn without replacement
    counts = rng.binomial(n, cfg.coverage, size=m)
    nnz = int(counts.sum())
    snv_col = np.repeat(np.arange(m, dtype=np.int64), counts)
    cell_col = np.empty(nnz, dtype=np.int64)
    # vectorised sampling without replacement: random keys, argpartition per SNV
    # is too slow at 1e8; instead sample with replacement then repair duplicates.
    cell_col[:] = rng.integers(0, n, size=nnz)
    key = snv_col * n + cell_col
    key = np.unique(key)                        # sorts by (snv, cell), drops dups
    snv_col = key // n
    cell_col = key % n
    nnz = key.shape[0]

    total = 1 + rng.poisson(cfg.extra_depth, size=nnz)
    present = anc[snv_cluster[snv_col], cell_clone[cell_col]]
    alt = rng.binomial(total, np.where(present, p_pres, p_abs))

    return SynthData(
        config=cfg,
        snv=snv_col.astype(np.int64),
        cell=cell_col.astype(np.int64),
        alt=alt.astype(np.int64),
        total=total.astype(np.int64),
        clust_snv=np.arange(m, dtype=np.int64),
        clust_label=noisy.astype(np.int64),
        trees=trees,
        true_tree_index=true_idx,
        cell_clone=cell_clone,
        snv_cluster=snv_cluster,
    )


def legacy_inputs(name="tiny"):
    """Inputs of the legacy scorer (``arborist_base.rank_trees``) from a synthetic workload: trees with the normal
    clone, read counts with the ``cluster`` column the reference's ``main`` derives from the SNV clustering
    (arborist_base.py:371-385).  One label is replaced by a cluster no tree contains, so the scorer's cluster
    filter (arborist_base.py:153-155) has rows to drop."""
    import importlib

    d = generate(CONFIGS[name])
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    rc = rc.copy()
    rc["cluster"] = rc["snv"].map(dict(zip(cl["snv"], cl["cluster"])))
    rc.loc[rc["snv"] % 17 == 5, "cluster"] = 99
    trees = importlib.import_module("arborist_b200.arborist").add_normal_clone(d.trees)
    return [[(int(u), int(v)) for u, v in t] for t in trees], rc
