"""``TreeFit`` -- the result type of one per-tree fit.

Same fields, methods and column names as the reference's dataclass
(``/root/reference/src/arborist/treefit.py:9-285``) so that callers, the CLI
writers and pickles keep working.  Two differences, both invisible to callers:

* ``q_z`` / ``q_y`` may be *lazy*: at the headline sizes a dict of 1,000-10,000
  eager fits is 2.7-236 GB (SURVEY.md section 7 "Output volume"), so
  ``TreeFit.lazy(...)`` keeps the posteriors on the GPU until first touched.
  Pickling always materialises them, so a pickle holds plain arrays.
* ``pygraphviz`` is imported inside ``visualize_tree`` only.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import pandas as pd

__all__ = ["TreeFit"]

_LAZY = ("q_z", "q_y")
_TRANSIENT = ("_loader", "_map_loader", "_tree_loader")


@dataclass
class TreeFit:
    tree: list
    tree_idx: int
    elbo: float
    q_z: np.ndarray
    q_y: np.ndarray
    cell_to_idx: dict
    snv_to_idx: dict
    clone_to_idx: dict
    cluster_to_idx: dict

    def __post_init__(self):
        self.idx_to_clone = {i: c for c, i in self.clone_to_idx.items()}
        self.idx_to_cluster = {i: c for c, i in self.cluster_to_idx.items()}

    # ---- lazy posteriors -------------------------------------------------
    @classmethod
    def lazy(cls, tree, tree_idx, elbo, loader, cell_to_idx, snv_to_idx, clone_to_idx,
             cluster_to_idx, map_loader=None, tree_loader=None):
        """Build a fit whose ``q_z``/``q_y`` are fetched by ``loader(tree_idx)`` on
        first access (returns ``(q_z, q_y)``); ``map_loader(tree_idx)`` optionally
        returns device-computed argmax vectors so MAP tables need no posterior copy;
        with ``tree=None`` the edge list comes from ``tree_loader(tree_idx)`` on first
        access (the ``--add-normal`` edge order is only built for trees that are looked at)."""
        self = object.__new__(cls)
        d = self.__dict__
        if tree is not None:
            d["tree"] = tree
        d.update(tree_idx=tree_idx, elbo=elbo, cell_to_idx=cell_to_idx,
                 snv_to_idx=snv_to_idx, clone_to_idx=clone_to_idx,
                 cluster_to_idx=cluster_to_idx, _loader=loader, _map_loader=map_loader,
                 _tree_loader=tree_loader)
        self.__post_init__()
        return self

    def __getattr__(self, name):
        # only reached when normal lookup fails, i.e. for not-yet-loaded posteriors / edge lists
        if name == "tree":
            tl = self.__dict__.get("_tree_loader")
            if tl is None:
                raise AttributeError(name)
            self.__dict__["tree"] = tl(self.tree_idx)
            return self.__dict__["tree"]
        if name in _LAZY:
            loader = self.__dict__.get("_loader")
            if loader is None:
                raise AttributeError(name)
            qz, qy = loader(self.tree_idx)
            self.__dict__["q_z"], self.__dict__["q_y"] = qz, qy
            return self.__dict__[name]
        raise AttributeError(name)

    def __getstate__(self):
        self.q_z  # noqa: B018 - force materialisation
        d = self.__dict__
        # the reference's field order, so the pickle reads like one of its own
        first = {k: (self.tree if k == "tree" else d[k])
                 for k in ("tree", "tree_idx", "elbo", "q_z", "q_y", "cell_to_idx", "snv_to_idx", "clone_to_idx",
                           "cluster_to_idx")}
        first.update({k: v for k, v in d.items() if k not in first and k not in _TRANSIENT})
        return first

    def __setstate__(self, state):
        self.__dict__.update(state)

    # ---- text ---------------------------------------------------------------
    def __str__(self):
        lines = [f"Tree {self.tree_idx}", f"ELBO: {self.elbo:.2f}"]
        lines += [f" {u}->{v}" for u, v in self.tree]
        return "\n".join(lines) + "\n"

    def __repr__(self):
        return (f"TreeFit(tree={self.tree}, tree_idx={self.tree_idx}, "
                f"expected_log_likelihood={self.elbo})")

    # ---- tables -------------------------------------------------------------
    @staticmethod
    def _row_labels(row_dict, n):
        inv = {i: k for k, i in row_dict.items()}
        return pd.Series(range(n)).map(inv)

    def _convert_q_to_dataframe(self, q, row_dict, col_dict, prefix="clone"):
        df = pd.DataFrame(q, columns=[f"{prefix}_{col_dict[i]}" for i in range(q.shape[1])])
        df.insert(0, "id", self._row_labels(row_dict, q.shape[0]))
        return df

    def q_z_df(self):
        """Cell-by-clone posterior as a DataFrame (``id``, ``clone_<label>``...)."""
        return self._convert_q_to_dataframe(self.q_z, self.cell_to_idx, self.idx_to_clone, "clone")

    def q_y_df(self):
        """SNV-by-cluster posterior as a DataFrame (``id``, ``cluster_<label>``...)."""
        return self._convert_q_to_dataframe(self.q_y, self.snv_to_idx, self.idx_to_cluster,
                                            "cluster")

    @staticmethod
    def _map_assign(q, mydict, assign_dict):
        """Row-wise argmax (first maximum wins, like ``ndarray.argmax``)."""
        return TreeFit._map_table(np.asarray(q).argmax(axis=1), mydict, assign_dict)

    @staticmethod
    def _map_table(arg, mydict, assign_dict):
        out = pd.DataFrame({"id": TreeFit._row_labels(mydict, len(arg)),
                            "assignment": [assign_dict[int(a)] for a in arg]})
        return out

    def _device_map(self):
        ml = self.__dict__.get("_map_loader")
        if ml is None or "q_z" in self.__dict__:
            return None
        return ml(self.tree_idx)

    def map_assign_z(self):
        """MAP clone of every cell: DataFrame ``['id', 'assignment']``."""
        dm = self._device_map()
        if dm is not None:
            return self._map_table(dm[0], self.cell_to_idx, self.idx_to_clone)
        return self._map_assign(self.q_z, self.cell_to_idx, self.idx_to_clone)

    def map_assign_y(self):
        """MAP cluster of every SNV: DataFrame ``['id', 'assignment']``."""
        dm = self._device_map()
        if dm is not None:
            return self._map_table(dm[1], self.snv_to_idx, self.idx_to_cluster)
        return self._map_assign(self.q_y, self.snv_to_idx, self.idx_to_cluster)

    # ---- files ----------------------------------------------------------------
    def save_tree(self, fname, sep=" "):
        """Write the tree in the input file's own block format."""
        with open(fname, "w+") as fh:
            fh.write(f"{len(self.tree)} #edges tree 0\n")
            fh.writelines(f"{u}{sep}{v}\n" for u, v in self.tree)

    def visualize_tree(self, output_file):
        """Graphviz drawing (``.dot`` written as text, anything else via ``dot``)."""
        from .utils import visualize_tree

        visualize_tree(self.tree, output_file=output_file)

    def _save_genotypes(self, fname, x=1, y=1):
        """Clone genotypes table (treefit.py:197-226); default copy numbers (1,1)."""
        snv_cluster = self.map_assign_y()
        parent = {v: u for u, v in self.tree}
        nodes = []
        for u, v in self.tree:
            for n in (u, v):
                if n not in nodes:
                    nodes.append(n)
        root = [n for n in nodes if n not in parent][0]

        def carries(node, cluster):
            # the reference tests identity or a DIRECT parent edge (has_predecessor)
            if cluster == root:
                return 0
            return int(node == cluster or parent.get(node) == cluster)

        rows = [[n, j, x, y, carries(n, c), 0, 1]
                for n in nodes for j, c in zip(snv_cluster["id"], snv_cluster["assignment"])]
        pd.DataFrame(rows, columns=["node", "snv", "x", "y", "xbar", "ybar", "segment"]).to_csv(
            fname, index=False)

    # ---- entropies ---------------------------------------------------------------
    def _compute_hz(self, eps=1e-12):
        return -np.sum(self.q_z * np.log(self.q_z + eps), axis=1) / self.q_z.shape[1]

    def _compute_hy(self, eps=1e-12):
        return -np.sum(self.q_y * np.log(self.q_y + eps), axis=1) / self.q_y.shape[1]

    def _entropy_df(self, h, row_dict):
        return pd.DataFrame({"id": self._row_labels(row_dict, h.shape[0]), "entropy": h})

    def cell_entropy(self, eps=1e-12):
        """Per-cell posterior entropy (divided by the number of clones, as upstream)."""
        return self._entropy_df(self._compute_hz(eps), self.cell_to_idx)

    def snv_cluster_entropy(self, eps=1e-12):
        """Per-SNV posterior entropy (divided by the number of clusters, as upstream)."""
        h = self._compute_hy(eps)
        assert h.shape[0] == len(self.snv_to_idx)
        return self._entropy_df(h, self.snv_to_idx)


# Pickles written by the reference name ``arborist.treefit.TreeFit``; the repo's
# ``arborist`` shim package re-exports this class under that path, and pickles
# written here use the same path so either package can read them.
TreeFit.__module__ = "arborist.treefit"
