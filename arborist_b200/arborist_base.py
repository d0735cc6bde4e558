"""Legacy tree scorer of the reference (``arborist.arborist_base``) on the GPU.

Mirrors /root/reference/src/arborist/arborist_base.py: ``build_genotypes`` (:228-247),
``find_cell_assignments`` (:102-194) and ``rank_trees`` (:250-362) with the same arguments and
return values.  The numeric work (binomial log-pmf of every read-count row, the per (cell, clone)
sums, best clone per cell, tree score, per-cell posteriors and entropies) runs in
``csrc/arb_base.cu`` through ``arb_base_ingest`` / ``arb_base_score`` / ``arb_base_cells``; this
module only factorises labels, builds the ancestor bit masks and assembles the DataFrames.  There is
no CPU path: without the CUDA library every entry point raises.

Difference from the reference, by design: all trees of one ``rank_trees`` call must have the same
node set (the reference concatenates frames with differing clone columns into NaNs and its
posterior step then uses the LAST tree's columns, arborist_base.py:303, 335-337).
"""
from __future__ import annotations

import numpy as np
import pandas as pd
from scipy.special import logsumexp

from .engine import Engine


def _clone_paths(tree):
    """(clones in the iteration order of the set of node labels, root-to-clone path of each)."""
    child_to_parent = {child: parent for parent, child in tree}
    all_clones = set(child_to_parent.keys()).union(set(child_to_parent.values()))
    clones, paths = [], []
    for clone in all_clones:
        path, x = [clone], clone
        while x in child_to_parent:
            x = child_to_parent[x]
            path.append(x)
        clones.append(clone)
        paths.append(path[::-1])
    return clones, paths


def build_genotypes(tree):
    """DataFrame ``Child`` / ``Ancestors`` ("root,...,child"), clones in the iteration order of the
    set of node labels (arborist_base.py:228-247)."""
    clones, paths = _clone_paths(tree)
    return pd.DataFrame([(c, ",".join(str(v) for v in p)) for c, p in zip(clones, paths)],
                        columns=["Child", "Ancestors"])


class _Table:
    """The read-count table on the device: rows with a cluster label folded per (cell, cluster)."""

    def __init__(self, read_counts, alpha, device=0):
        for col in ("cell", "cluster", "total", "alt"):
            if col not in read_counts.columns:
                raise ValueError(f"read_counts needs a '{col}' column")
        rc = read_counts[read_counts["cluster"].notna()]
        if len(rc) == 0:
            raise ValueError("read_counts has no row with a cluster label")
        self.cell_code, self.cell_ids = pd.factorize(rc["cell"], sort=False)
        self.cluster_code, cluster_ids = pd.factorize(rc["cluster"].astype(np.int64), sort=True)
        self.cluster_ids = [int(c) for c in cluster_ids]
        self._bit = {c: 1 << k for k, c in enumerate(self.cluster_ids)}
        self.engine = Engine(device)
        self.engine.base_ingest(self.cell_code, self.cluster_code, rc["total"].to_numpy(), rc["alt"].to_numpy(),
                                len(self.cell_ids), len(self.cluster_ids), alpha)

    def masks(self, genotype_matrix):
        """(ancestor mask per clone, mask of the tree's own clusters) over the cluster codes."""
        clones = [int(c) for c in genotype_matrix["Child"].values]
        paths = [[int(x) for x in str(p).split(",")] for p in genotype_matrix["Ancestors"].values]
        return self.path_masks(clones, paths)

    def path_masks(self, clones, paths):
        bit = self._bit
        anc = np.zeros(len(clones), np.uint32)
        for j, path in enumerate(paths):
            m = 0
            for c in path:
                m |= bit.get(int(c), 0)
            anc[j] = m
        tmask = 0
        for c in clones:
            tmask |= bit.get(int(c), 0)
        return anc, np.uint32(tmask)

    def cell_order(self, tmask):
        """Cells that keep a row under the tree's cluster filter, in first-appearance order of those rows
        (arborist_base.py:153-160)."""
        keep = ((int(tmask) >> self.cluster_code) & 1).astype(bool)
        return pd.unique(self.cell_code[keep])

    def close(self):
        self.engine.close()


def find_cell_assignments(read_counts, genotype_matrix, error_rate=0.001, device=0):
    """``(score, Cell_assignment)``: log-likelihood of every cell (rows) under every clone (columns) and the
    sum over cells of the row maxima (arborist_base.py:102-194)."""
    tab = _Table(read_counts, error_rate, device)
    try:
        anc, tmask = tab.masks(genotype_matrix)
        score = float(tab.engine.base_score(anc[None, :], np.array([tmask], np.uint32))[0])
        ll, _, _, _, _ = tab.engine.base_cells(anc[None, :], np.array([tmask], np.uint32), want_loglik=True)
        order = tab.cell_order(tmask)
        frame = pd.DataFrame(ll[0][order], index=list(tab.cell_ids[order]), columns=genotype_matrix["Child"].values,
                             dtype=float)
        return score, frame
    finally:
        tab.close()


def rank_trees(tree_list, read_counts, alpha=0.001, topn=None, verbose=False, device=0):
    """``(ranked_trees, cell_assignments)`` like the reference (arborist_base.py:250-362): trees scored by the sum
    over cells of the best clone's log-likelihood, ``posterior`` in percent over all trees, then for the ``topn``
    best trees one row per cell with its best clone, the entropy (bits) of its clone posterior and the posterior
    itself."""
    if len(tree_list) == 0:
        raise ValueError("rank_trees: empty tree list")
    shapes = [_clone_paths(tree) for tree in tree_list]
    orders = [[int(c) for c in clones] for clones, _ in shapes]
    node_set = sorted(orders[0])
    for t, o in enumerate(orders):
        if sorted(o) != node_set:
            raise ValueError(f"rank_trees: tree {t} has a different node set than tree 0")
    tab = _Table(read_counts, alpha, device)
    try:
        pairs = [tab.path_masks(clones, paths) for clones, paths in shapes]
        anc = np.stack([p[0] for p in pairs])
        tmask = np.array([p[1] for p in pairs], np.uint32)
        lik = tab.engine.base_score(anc, tmask)
        if verbose:
            for t, v in enumerate(lik):
                print(f"Processed tree {t} with likelihood {v}.")
            print(f"Processed {len(tree_list)} trees.")
        df = pd.DataFrame({"tree": np.arange(len(tree_list)), "likelihood": lik})
        df["posterior"] = np.exp(df["likelihood"] - logsumexp(df["likelihood"])) * 100
        df.sort_values(by=["likelihood"], ascending=False, inplace=True)
        df.reset_index(drop=True, inplace=True)
        ranked_trees = df.iloc[:topn]

        sel = sorted(int(t) for t in ranked_trees["tree"])
        _, post, idx, ent, _ = tab.engine.base_cells(anc[sel], tmask[sel])
        # the reference names the posterior columns after the LAST tree's clone order (arborist_base.py:303)
        last = orders[-1]
        frames = []
        order = tab.cell_order(tmask[0])                  # same node set, so the same cluster filter for every tree
        for s, t in enumerate(sel):
            mine = orders[t]
            col = [mine.index(c) for c in last]
            f = pd.DataFrame({"cell": tab.cell_ids[order], "tree": t, "clone": np.asarray(mine)[idx[s][order]],
                              "entropy": ent[s][order]})
            for j, c in zip(col, last):
                f[f"clone_{c}_posterior"] = post[s][order, j]
            # row labels of the reference: positions in the concatenation over ALL trees (arborist_base.py:327-331)
            f.index = np.arange(len(order)) + t * len(order)
            frames.append(f)
        cell_assignments = pd.concat(frames)
        return ranked_trees, cell_assignments
    finally:
        tab.close()


# ---------------------------------------------------------------------------
# writers and CLI (arborist_base.py:13-99, 197-225, 365-424) -- host only
# ---------------------------------------------------------------------------
def save_tree(fname, tree, sep=" "):
    """Edge list with the reference's one-line header (arborist_base.py:94-99)."""
    with open(fname, "w+") as f:
        f.write(f"{len(tree)} #edges tree 0\n")
        for u, v in tree:
            f.write(f"{u}{sep}{v}\n")


def save_genotypes(fname, tree, snv_to_cluster, x=1, y=1):
    """Long genotype table node x SNV (arborist_base.py:197-225).  Like the reference, a node carries a cluster's
    SNVs when the cluster is the node itself or its PARENT (``has_predecessor`` is the direct-edge test), never the
    root's; SNVs of clusters outside the tree get 0."""
    from .arborist import _digraph_edge_order

    edges, nodes = _digraph_edge_order(tree)
    parents = {}
    for u, v in edges:
        parents.setdefault(v, set()).add(u)
    root = [n for n in nodes if not parents.get(n)][0]
    node_set, rows = set(nodes), []
    for n in nodes:
        for snv, clust in snv_to_cluster.items():
            xbar = int(clust in node_set and clust != root and (n == clust or clust in parents.get(n, ())))
            rows.append([n, snv, x, y, xbar, 0, 1])
    pd.DataFrame(rows, columns=["node", "snv", "x", "y", "xbar", "ybar", "segment"]).to_csv(fname, index=False)


def parse_arguments(argv=None):
    import argparse

    p = argparse.ArgumentParser(description="Tree ranking script")
    p.add_argument("-R", "--read_counts", required=True,
                   help="Path to read counts CSV file with columns 'cell', 'total', 'alt'")
    p.add_argument("-C", "--clustering", required=True,
                   help="path to snv cluster labels, csv file with unlabeled columns 'snv', 'cluster'")
    p.add_argument("-T", "--trees", required=True, help="Path to tree file")
    p.add_argument("--edge-delim", type=str, default=" ", help="edge delimiter in candidate tree file.")
    p.add_argument("--alpha", type=float, default=0.001, help="Per base sequencing error")
    p.add_argument("--topn", type=int, default=25, help="Filter only the top n trees, default is 25.")
    p.add_argument("--ranking", type=str, help="Path to where tree ranking output should be saved")
    p.add_argument("--cell-assign", type=str, help="Path to where cell assignments output should be saved")
    p.add_argument("-v", "--verbose", help="Print verbose output", action="store_true")
    p.add_argument("-d", "--draw", type=str, help="Path to save the tree image")
    p.add_argument("--snv-assign", type=str, help="Path to where snv assignments output should be saved")
    p.add_argument("--genotypes", type=str, help="Path to where the genotypes should be saved")
    p.add_argument("--add-normal", action="store_true",
                   help="indicates if a normal root should be added to the tree")
    p.add_argument("-t", "--tree", type=str, help="path to where the tree should be saved")
    return p.parse_args(argv)


def main(argv=None):
    """The reference's legacy command line (arborist_base.py:365-424)."""
    from .arborist import add_normal_clone
    from .utils import read_read_counts, read_trees, visualize_tree

    args = parse_arguments(argv)
    read_counts = read_read_counts(args.read_counts)
    clustering = pd.read_csv(args.clustering, header=None, names=["snv", "cluster"])
    snv_to_cluster = dict(zip(clustering["snv"], clustering["cluster"]))
    read_counts["cluster"] = read_counts["snv"].map(snv_to_cluster)
    candidate_trees = read_trees(args.trees, sep=args.edge_delim)
    if args.add_normal:
        candidate_trees = add_normal_clone(candidate_trees)
    ranked_trees, cell_assignments = rank_trees(candidate_trees, read_counts, alpha=args.alpha, topn=args.topn,
                                                verbose=args.verbose)
    tree_idx = ranked_trees["tree"].tolist()[0]
    best_tree = cell_assignments[cell_assignments["tree"] == tree_idx]
    inf_tree = candidate_trees[tree_idx]
    if args.draw:
        visualize_tree(inf_tree, cell_assignment=dict(zip(best_tree["cell"], best_tree["clone"])),
                       output_file=args.draw)
    if args.tree:
        save_tree(args.tree, inf_tree)
    if args.genotypes:
        save_genotypes(args.genotypes, inf_tree, snv_to_cluster)
    if args.snv_assign:
        nodes = {x for e in inf_tree for x in e}
        kept = clustering[clustering["cluster"].isin(list(nodes))]
        alt_sum = read_counts.groupby("snv")["alt"].sum()
        valid_snvs = alt_sum[alt_sum > 0].index
        print(f"Number of valid SNVs:  {len(valid_snvs)}")
        kept[kept["snv"].isin(valid_snvs)].to_csv(args.snv_assign, index=False)
    if args.ranking:
        ranked_trees.to_csv(args.ranking, index=False)
    if args.cell_assign:
        best_tree[["cell", "clone"]].copy().to_csv(args.cell_assign, index=False)


if __name__ == "__main__":
    main()
