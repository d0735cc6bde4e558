"""CPU restatement of the reference's legacy tree scorer -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the
product path (arborist_b200.arborist_base) never does.

Follows /root/reference/src/arborist/arborist_base.py:
  * build_genotypes        (arborist_base.py:228-247): clone -> its ancestors (itself included), clones in
                           the iteration order of the Python set of all node labels;
  * find_cell_assignments  (arborist_base.py:102-194): rows whose cluster is not a node of the tree are dropped;
                           cells in first-appearance order of the remaining rows; for every (cell, clone) the sum
                           over the cell's rows of binom.logpmf(alt, total, p) with p = 0.5 - e + 0.5 e / 3 if the
                           row's cluster is an ancestor of the clone, e / 3 otherwise; the tree's score is the sum
                           over cells of the best clone's value;
  * rank_trees             (arborist_base.py:250-362): scores -> posterior (percent) over trees, descending sort,
                           top-n filter, per-cell posteriors over clones (softmax of the log-likelihoods), MAP clone
                           (first maximum in column order) and entropy in bits.

Pinned by tests/golden/base_tiny.npz, written by oracle/gen_base_golden.py from the live reference.

The restatement aggregates per (cell, cluster) first -- the per-row log-pmf does not depend on the tree -- so it is a
vectorised numpy program, not the reference's iterrows loop; sums differ from the reference's single np.sum by
rounding only (~1e-15 relative).
"""
from __future__ import annotations

import numpy as np
import pandas as pd
from scipy.special import logsumexp
from scipy.stats import binom


def clone_order(tree):
    """Node labels in the order build_genotypes lists them (arborist_base.py:229-230, 240-244)."""
    child_to_parent = {c: p for p, c in tree}
    return list(set(child_to_parent.keys()).union(set(child_to_parent.values()))), child_to_parent


def ancestor_sets(tree):
    clones, c2p = clone_order(tree)
    out = []
    for c in clones:
        anc, x = {int(c)}, c
        while x in c2p:
            x = c2p[x]
            anc.add(int(x))
        out.append(anc)
    return clones, out


def cell_tables(read_counts, alpha):
    """Tree-independent part: per (cell, cluster) sums of log-pmf at p_mutated / p_not_mutated and row counts."""
    rc = read_counts[read_counts["cluster"].notna()]
    cells, cell_ids = pd.factorize(rc["cell"], sort=False)
    clus, clus_ids = pd.factorize(rc["cluster"].astype(np.int64), sort=True)
    a, d = rc["alt"].to_numpy(np.int64), rc["total"].to_numpy(np.int64)
    p_mut = 0.5 - alpha + 0.5 * alpha / 3            # arborist_base.py:165
    p_not = alpha / 3                                # arborist_base.py:166
    n, C = len(cell_ids), len(clus_ids)
    TA, TB, CNT = np.zeros((n, C)), np.zeros((n, C)), np.zeros((n, C), np.int64)
    np.add.at(TA, (cells, clus), binom.logpmf(a, d, p_mut))
    np.add.at(TB, (cells, clus), binom.logpmf(a, d, p_not))
    np.add.at(CNT, (cells, clus), 1)
    first = np.full(n, np.iinfo(np.int64).max)
    np.minimum.at(first, cells, np.arange(len(cells)))
    return dict(TA=TA, TB=TB, CNT=CNT, cell_ids=np.asarray(cell_ids), cluster_ids=np.asarray(clus_ids, np.int64),
                cells=cells, clus=clus)


def score_tree(tab, tree):
    """(score, cells kept [indices into tab['cell_ids'] in first-appearance order], clones, L [cells, clones])."""
    clones, anc = ancestor_sets(tree)
    in_tree = np.isin(tab["cluster_ids"], np.asarray([int(c) for c in clones], np.int64))
    # cells in first-appearance order of the rows that survive the cluster filter (arborist_base.py:153-160)
    keep_row = in_tree[tab["clus"]]
    kept_cells = pd.unique(tab["cells"][keep_row])
    L = np.zeros((len(kept_cells), len(clones)))
    for j, s in enumerate(anc):
        is_anc = np.isin(tab["cluster_ids"], np.asarray(sorted(s), np.int64))
        use = np.where(is_anc[None, :], tab["TA"][kept_cells], tab["TB"][kept_cells])
        L[:, j] = (use * in_tree[None, :]).sum(axis=1)
    return float(L.max(axis=1).sum()), kept_cells, clones, L


def rank_trees_oracle(tree_list, read_counts, alpha=0.001, topn=None):
    tab = cell_tables(read_counts, alpha)
    scored = [score_tree(tab, t) for t in tree_list]
    lik = np.array([s[0] for s in scored])
    df = pd.DataFrame({"tree": np.arange(len(tree_list)), "likelihood": lik})
    df["posterior"] = np.exp(df["likelihood"] - logsumexp(df["likelihood"])) * 100     # arborist_base.py:318-319
    df.sort_values(by=["likelihood"], ascending=False, inplace=True)
    df.reset_index(drop=True, inplace=True)
    ranked = df.iloc[:topn]
    rows = []
    for t in sorted(ranked["tree"].tolist()):
        _, kept, clones, L = scored[t]
        post = np.exp(L - logsumexp(L, axis=1, keepdims=True))                         # arborist_base.py:335-338
        pn = post / post.sum(axis=1, keepdims=True)
        with np.errstate(divide="ignore", invalid="ignore"):
            ent = -np.where(pn > 0, pn * np.log(pn), 0.0).sum(axis=1) / np.log(2.0)    # scipy entropy(base=2)
        rows.append(dict(tree=t, cells=tab["cell_ids"][kept], clone=np.asarray(clones)[L.argmax(axis=1)],
                         entropy=ent, posterior=post, clones=clones, loglik=L))
    return ranked, rows
