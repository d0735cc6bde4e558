"""Legacy scorer (reference arborist_base.py): the numpy oracle against the golden recorded from the unmodified
reference (CPU), and the CUDA path against both (GPU).  Scores / log-likelihoods within 1e-9 relative (observed
~1e-15), rankings, best clones and cell orders exact."""
import numpy as np
import pytest

from conftest import load_golden

TOPN = 3
RTOL = 1e-9


def _inputs():
    from arborist_b200 import synth
    return synth.legacy_inputs("tiny")


def test_oracle_matches_reference_golden():
    from oracle import oracle_base as ob
    g = load_golden("base_tiny")
    trees, rc = _inputs()
    ranked, rows = ob.rank_trees_oracle(trees, rc, 0.001, TOPN)
    assert ranked["tree"].tolist() == g["ranked_tree"].tolist()
    np.testing.assert_allclose(ranked["likelihood"], g["ranked_likelihood"], rtol=1e-12)
    np.testing.assert_allclose(ranked["posterior"], g["ranked_posterior"], rtol=1e-9)
    tab = ob.cell_tables(rc, 0.001)
    for t, tree in enumerate(trees):
        score, kept, clones, L = ob.score_tree(tab, tree)
        assert list(tab["cell_ids"][kept]) == list(g["table_cells"][t])      # cells dropped by the cluster filter too
        assert list(clones) == list(g["table_clones"][t])
        assert score == pytest.approx(float(g["score"][t]), rel=1e-12)
        np.testing.assert_allclose(L, g["table_loglik"][t], rtol=1e-12, atol=1e-12)
    for r in rows:
        sel = g["cells_tree"] == r["tree"]
        assert (g["cells_cell"][sel] == r["cells"]).all() and (g["cells_clone"][sel] == r["clone"]).all()
        np.testing.assert_allclose(r["entropy"], g["cells_entropy"][sel], rtol=1e-9, atol=1e-13)
        np.testing.assert_allclose(r["posterior"], g["cells_posterior"][sel], rtol=1e-9, atol=1e-15)


def test_build_genotypes_matches_reference_golden():
    from arborist_b200 import arborist_base as ab
    g = load_golden("base_tiny")
    trees, _ = _inputs()
    for t, tree in enumerate(trees):
        gm = ab.build_genotypes(tree)
        assert [int(c) for c in gm["Child"]] == list(g["table_clones"][t])
        for child, path in zip(gm["Child"], gm["Ancestors"]):
            nodes = [int(x) for x in path.split(",")]
            assert nodes[-1] == child and nodes[0] == 0                      # root-to-clone path, normal clone first
            assert all((u, v) in tree for u, v in zip(nodes, nodes[1:]))


@pytest.mark.gpu
def test_rank_trees_matches_reference_golden():
    from arborist_b200 import arborist_base as ab
    g = load_golden("base_tiny")
    trees, rc = _inputs()
    ranked, cells = ab.rank_trees(trees, rc, alpha=0.001, topn=TOPN)
    assert list(ranked.columns) == ["tree", "likelihood", "posterior"]
    assert ranked["tree"].tolist() == g["ranked_tree"].tolist()
    np.testing.assert_allclose(ranked["likelihood"], g["ranked_likelihood"], rtol=RTOL)
    np.testing.assert_allclose(ranked["posterior"], g["ranked_posterior"], rtol=1e-7)
    assert list(cells.columns) == ["cell", "tree", "clone", "entropy", *g["posterior_columns"].tolist()]
    assert cells.index.tolist() == g["cells_index"].tolist()
    assert cells["cell"].tolist() == g["cells_cell"].tolist() and cells["tree"].tolist() == g["cells_tree"].tolist()
    assert cells["clone"].tolist() == g["cells_clone"].tolist()
    np.testing.assert_allclose(cells["entropy"], g["cells_entropy"], rtol=1e-7, atol=1e-12)
    np.testing.assert_allclose(cells[g["posterior_columns"].tolist()].to_numpy(), g["cells_posterior"], rtol=1e-7,
                               atol=1e-14)


@pytest.mark.gpu
def test_find_cell_assignments_matches_reference_golden():
    from arborist_b200 import arborist_base as ab
    g = load_golden("base_tiny")
    trees, rc = _inputs()
    for t in (0, 3, 5):
        score, frame = ab.find_cell_assignments(rc, ab.build_genotypes(trees[t]), 0.001)
        assert score == pytest.approx(float(g["score"][t]), rel=RTOL)
        assert frame.index.tolist() == g["table_cells"][t].tolist()
        assert [int(c) for c in frame.columns] == g["table_clones"][t].tolist()
        np.testing.assert_allclose(frame.to_numpy(), g["table_loglik"][t], rtol=RTOL, atol=1e-10)


@pytest.mark.gpu
def test_rank_trees_against_oracle_on_a_larger_table():
    """config 2 (1,000 cells x 5,000 SNVs, 100 trees): scores, ranking, best clones vs the oracle;
    string cell ids; two runs give identical bits."""
    from arborist_b200 import arborist_base as ab, synth
    from oracle import oracle_base as ob
    trees, rc = synth.legacy_inputs("config2")
    rc = rc.copy()
    rc["cell"] = "c" + rc["cell"].astype(str)
    ranked, cells = ab.rank_trees(trees, rc, alpha=0.002, topn=5)
    oranked, orows = ob.rank_trees_oracle(trees, rc, 0.002, 5)
    np.testing.assert_allclose(ranked.sort_values("tree")["likelihood"], oranked.sort_values("tree")["likelihood"],
                               rtol=RTOL)
    assert ranked["tree"].tolist() == oranked["tree"].tolist()
    for r in orows:
        mine = cells[cells["tree"] == r["tree"]]
        assert mine["cell"].tolist() == list(r["cells"])
        diff = np.nonzero(mine["clone"].to_numpy() != r["clone"])[0]
        for i in diff:                                   # only where the oracle's own maximum is a rounding tie
            L = r["loglik"][i]
            assert abs(L[list(r["clones"]).index(mine["clone"].iloc[i])] - L.max()) <= 1e-9 * abs(L.max())
        np.testing.assert_allclose(mine["entropy"], r["entropy"], rtol=1e-6, atol=1e-10)
    ranked2, cells2 = ab.rank_trees(trees, rc, alpha=0.002, topn=5)
    assert (ranked2["likelihood"].to_numpy() == ranked["likelihood"].to_numpy()).all()
    assert (cells2["entropy"].to_numpy() == cells["entropy"].to_numpy()).all()


@pytest.mark.gpu
def test_rank_trees_errors():
    from arborist_b200 import arborist_base as ab
    trees, rc = _inputs()
    with pytest.raises(ValueError):
        ab.rank_trees([], rc)
    with pytest.raises(ValueError):
        ab.rank_trees([trees[0], [(0, 1), (1, 2)]], rc)          # different node sets
    with pytest.raises(ValueError):
        ab.rank_trees(trees, rc.drop(columns=["cluster"]))


@pytest.mark.gpu
def test_legacy_cli_outputs_match_reference(tmp_path):
    """The reference's legacy command line (arborist_base.py:365-424) replayed on the files of
    tests/golden/cli_tiny; expected outputs written by oracle/gen_base_golden.py from the live reference."""
    import os
    import pandas as pd
    from conftest import GOLDEN
    from arborist_b200 import arborist_base as ab
    src, exp = os.path.join(GOLDEN, "cli_tiny"), os.path.join(GOLDEN, "base_cli_tiny")
    outputs = {"--ranking": "ranking.csv", "--cell-assign": "cell_assign.csv", "--snv-assign": "snv_assign.csv",
               "--genotypes": "genotypes.csv", "-t": "tree.txt"}
    argv = ["-R", os.path.join(src, "read_counts.csv"), "-C", os.path.join(src, "clusters.csv"),
            "-T", os.path.join(src, "trees.txt"), "--add-normal", "--alpha", "0.001", "--topn", "2"]
    for flag, name in outputs.items():
        argv += [flag, str(tmp_path / name)]
    ab.main(argv)
    for name in ("cell_assign.csv", "snv_assign.csv", "genotypes.csv", "tree.txt"):
        assert (tmp_path / name).read_text() == open(os.path.join(exp, name)).read(), name
    got, want = pd.read_csv(tmp_path / "ranking.csv"), pd.read_csv(os.path.join(exp, "ranking.csv"))
    assert list(got.columns) == list(want.columns) and got["tree"].tolist() == want["tree"].tolist()
    np.testing.assert_allclose(got["likelihood"], want["likelihood"], rtol=RTOL)
    np.testing.assert_allclose(got["posterior"], want["posterior"], rtol=1e-7)


def test_legacy_cli_flags_match_reference():
    """Same flags and defaults as arborist_base.py:13-91."""
    from arborist_b200 import arborist_base as ab
    a = ab.parse_arguments(["-R", "r", "-C", "c", "-T", "t"])
    assert (a.alpha, a.topn, a.edge_delim, a.add_normal, a.verbose) == (0.001, 25, " ", False, False)
    assert all(getattr(a, k) is None for k in ("ranking", "cell_assign", "snv_assign", "genotypes", "tree", "draw"))
