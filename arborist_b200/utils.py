"""Host-side helpers kept from the reference's surface.

``read_trees`` mirrors ``/root/reference/src/arborist/utils.py:73-141`` (same
file grammar, same return type).  Drawing needs ``pygraphviz`` + Graphviz's
``dot``; neither is in this image, so it is imported lazily and reported as
unavailable instead of breaking ``import arborist_b200``.
"""
from __future__ import annotations

__all__ = ["read_trees", "read_read_counts", "visualize_tree", "write_trees"]


def read_trees(file_path, sep=" "):
    """Parse a candidate-tree file into ``[[(parent, child), ...], ...]``.

    Grammar (utils.py:121-139): a line that is empty or contains ``#`` closes the
    current tree; every other line is split on ``sep`` and kept only when it has
    exactly two fields, both parsed as ``int``.
    """
    trees, cur = [], []
    with open(file_path, "r") as fh:
        for raw in fh:
            line = raw.strip()
            if not line or "#" in line:
                if cur:
                    trees.append(cur)
                    cur = []
                continue
            parts = line.split(sep)
            if len(parts) == 2:
                cur.append((int(parts[0]), int(parts[1])))
    if cur:
        trees.append(cur)
    return trees


def read_read_counts(path):
    """The long-format read-count table (columns ``snv, cell, alt, total``) as a DataFrame.

    The reference reads a CSV (``pd.read_csv``, A.py:1135), which at 10^8 rows takes minutes and
    becomes the floor of a run once the fit is on the GPU.  Besides CSV
    this accepts the same four columns as Parquet / Feather (pyarrow) or as an ``.npz`` of four
    arrays; integer id columns then go to the device as they are (``arb_ingest_ids``)."""
    import pandas as pd

    low = str(path).lower()
    if low.endswith((".parquet", ".pq")):
        df = pd.read_parquet(path)
    elif low.endswith((".feather", ".arrow")):
        df = pd.read_feather(path)
    elif low.endswith(".npz"):
        import numpy as np

        with np.load(path, allow_pickle=False) as z:
            df = pd.DataFrame({c: z[c] for c in ("snv", "cell", "alt", "total")})
    else:
        return pd.read_csv(path)
    missing = [c for c in ("snv", "cell", "alt", "total") if c not in df.columns]
    if missing:
        raise ValueError(f"{path}: missing column(s) {missing}")
    return df


def write_trees(trees, file_path, sep=" "):
    """Inverse of :func:`read_trees` (same block layout as ``TreeFit.save_tree``)."""
    with open(file_path, "w") as fh:
        for t, edges in enumerate(trees):
            fh.write(f"{len(edges)} #edges tree {t}\n")
            for u, v in edges:
                fh.write(f"{u}{sep}{v}\n")


def visualize_tree(tree, cell_assignment=None, output_file=None):
    """Graphviz rendering of an edge list (utils.py:145-194). Needs pygraphviz."""
    try:
        import pygraphviz as pgv
    except ImportError as exc:  # pragma: no cover - not installable in this image
        raise RuntimeError(
            "drawing needs pygraphviz and Graphviz 'dot' (unavailable on this box)") from exc
    counts = {}
    if cell_assignment is not None:
        for v in dict(cell_assignment).values():
            counts[v] = counts.get(v, 0) + 1
    g = pgv.AGraph(directed=True)
    for u, v in tree:
        for x in (u, v):
            label = f"{x}\n{counts[x]} cells" if x in counts else f"{x}"
            g.add_node(x, shape="circle", label=label)
        g.add_edge(u, v)
    if output_file:
        if output_file.endswith(".dot"):
            g.write(output_file)
        else:
            g.draw(output_file, prog="dot")
    return g
