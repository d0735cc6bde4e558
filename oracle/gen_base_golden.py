"""Generate tests/golden/base_tiny.npz from the UNMODIFIED reference's legacy scorer (build container only).

    python oracle/gen_base_golden.py

Runs ``arborist.arborist_base.rank_trees`` and ``find_cell_assignments`` (arborist_base.py:102-194, 250-362) of
/root/reference on the ``tiny`` synthetic workload: inputs from ``arborist_b200.synth.legacy_inputs`` (trees with the normal clone, read counts with
the ``cluster`` column of arborist_base.py:371-373, one label replaced by a cluster no tree contains so the
filter of arborist_base.py:153-155 is exercised).  Inputs are regenerated from the seed by the tests; only outputs are stored.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
OUT = os.path.join(ROOT, "tests", "golden", "base_tiny.npz")
TOPN = 3


def inputs():
    """(trees with the normal clone, read counts with a cluster column) of the tiny workload."""
    from arborist_b200 import synth

    return synth.legacy_inputs("tiny")


CLI_OUTPUTS = {"--ranking": "ranking.csv", "--cell-assign": "cell_assign.csv", "--snv-assign": "snv_assign.csv",
               "--genotypes": "genotypes.csv", "-t": "tree.txt"}


def cli_args(indir, outdir):
    argv = ["-R", os.path.join(indir, "read_counts.csv"), "-C", os.path.join(indir, "clusters.csv"),
            "-T", os.path.join(indir, "trees.txt"), "--add-normal", "--alpha", "0.001", "--topn", "2"]
    for flag, name in CLI_OUTPUTS.items():
        argv += [flag, os.path.join(outdir, name)]
    return argv


def main():
    from gen_golden import load_reference

    load_reference()
    import importlib

    B = importlib.import_module("arborist.arborist_base")
    trees, rc = inputs()
    with contextlib.redirect_stdout(io.StringIO()):      # the reference prints every tree and table head
        ranked, cells = B.rank_trees(trees, rc, alpha=0.001, topn=TOPN)
        tables = [B.find_cell_assignments(rc, B.build_genotypes(t), 0.001) for t in trees]
    post_cols = [c for c in cells.columns if c.endswith("_posterior")]
    np.savez_compressed(
        OUT,
        ranked_tree=ranked["tree"].to_numpy(np.int64), ranked_likelihood=ranked["likelihood"].to_numpy(np.float64),
        ranked_posterior=ranked["posterior"].to_numpy(np.float64),
        cells_index=cells.index.to_numpy(np.int64), cells_cell=cells["cell"].to_numpy(np.int64),
        cells_tree=cells["tree"].to_numpy(np.int64), cells_clone=cells["clone"].to_numpy(np.int64),
        cells_entropy=cells["entropy"].to_numpy(np.float64), cells_posterior=cells[post_cols].to_numpy(np.float64),
        posterior_columns=np.array(post_cols),
        score=np.array([t[0] for t in tables]),
        table_cells=np.stack([t[1].index.to_numpy(np.int64) for t in tables]),
        table_clones=np.stack([t[1].columns.to_numpy(np.int64) for t in tables]),
        table_loglik=np.stack([t[1].to_numpy(np.float64) for t in tables]),
    )
    print("wrote", OUT, "ranking", ranked["tree"].tolist(), "rows", len(cells))

    # the legacy command line (arborist_base.py:365-424) on the files of tests/golden/cli_tiny
    src = os.path.join(ROOT, "tests", "golden", "cli_tiny")
    exp = os.path.join(ROOT, "tests", "golden", "base_cli_tiny")
    os.makedirs(exp, exist_ok=True)
    old = sys.argv
    sys.argv = ["arborist_base", *cli_args(src, exp)]
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            B.main()
    finally:
        sys.argv = old
    print("wrote", sorted(os.listdir(exp)))


if __name__ == "__main__":
    main()
