"""Host-side logic that mirrors the reference's Python surface (no GPU needed)."""
import pickle

import numpy as np
import pandas as pd
import pytest

from conftest import golden_frames, load_golden


def test_read_trees_roundtrip(tmp_path):
    from arborist_b200.utils import read_trees, write_trees
    trees = [[(0, 1), (1, 2)], [(0, 2), (2, 1)], [(5, 6)]]
    p = tmp_path / "t.txt"
    write_trees(trees, p)
    assert read_trees(p) == trees
    p.write_text("0,1\n1,2\n# tree\n\n0,2\n2,1\nbad line here\n")
    assert read_trees(p, sep=",") == trees[:2]


def test_add_normal_clone_matches_reference_edge_order():
    """A.py:860-882 returns list(nx.DiGraph.edges) of the augmented graph."""
    from arborist_b200.arborist import add_normal_clone
    g = load_golden("small")
    given = [[(int(u), int(v)) for u, v in t] for t in g["trees"]]
    g2 = load_golden("small_nonormal")           # its trees = add_normal_clone(small's trees), via the reference path
    want = [[(int(u), int(v)) for u, v in t] for t in g2["trees"]]
    assert add_normal_clone(given) == want
    assert add_normal_clone([[(3, 4), (3, 5)]]) == [[(3, 4), (3, 5), (2, 3)]]


def test_prepare_trees_parents_and_errors():
    from arborist_b200.arborist import prepare_trees
    trees = [[(0, 2), (2, 5), (2, 7)], [(0, 5), (5, 2), (2, 7)]]
    c2i, k2i, parents, presence, normal = prepare_trees(trees)
    assert c2i == {0: 0, 2: 1, 5: 2, 7: 3} and k2i == {2: 0, 5: 1, 7: 2} and normal == 0
    assert presence is None
    assert parents.tolist() == [[-1, 0, 1, 1], [-1, 2, 0, 1]]
    with pytest.raises(ValueError, match="same set of clones"):
        prepare_trees([[(0, 1), (1, 2)], [(0, 1), (1, 3)]])
    with pytest.raises(ValueError, match="same set of clones"):
        prepare_trees([[(0, 1), (1, 2)], [(0, 1), (0, 1)]])
    # ragged edge lists take the slow path
    c2i, _, parents, presence, _ = prepare_trees([[(0, 1), (1, 2)], [(0, 1), (1, 2), (0, 1)]])
    assert presence is None and parents.tolist() == [[-1, 0, 1], [-1, 0, 1]]


def test_prepare_trees_non_forest_gives_presence():
    from arborist_b200.arborist import prepare_trees
    from oracle import oracle as orc
    dag = [[(0, 1), (1, 2), (1, 3), (2, 4), (3, 4)], [(0, 1), (1, 2), (2, 3), (3, 4)]]
    c2i, k2i, parents, presence, normal = prepare_trees(dag)
    assert parents is None and presence.shape == (2, 4, 5)
    for t in range(2):
        assert np.array_equal(presence[t], orc.enumerate_presence(dag[t], c2i, k2i))


@pytest.mark.parametrize("name", ["tiny", "deep30"])
def test_parent_closure_equals_reference_presence(name):
    """The closure the device builds from parent vectors, restated in numpy, equals
    enumerate_presence (A.py:491-520) as recorded from the reference."""
    from arborist_b200.arborist import add_normal_clone, prepare_trees
    g = load_golden(name)
    trees, _, _ = golden_frames(g)
    trees = add_normal_clone(trees)
    c2i, k2i, parents, _, normal = prepare_trees(trees)
    K = len(c2i)
    for t in range(len(trees)):
        P = np.zeros((K - 1, K), dtype=np.int64)
        for r in range(K):
            a = r
            while a >= 0:
                if a != normal:
                    P[a if a < normal else a - 1, r] = 1
                a = parents[t, a]
        assert np.array_equal(P, g["presence"][t])


def test_prior_labels_quirks():
    """A.py:547-567: last duplicate wins, label 0 / off-tree / missing -> uniform row."""
    from arborist_b200.arborist import _prior_labels
    cl = pd.DataFrame({"snv": [10, 11, 12, 13, 10], "cluster": [1, 0, 99, 2, 2]})
    out = _prior_labels(cl, np.array([10, 11, 12, 13, 14]), {1: 0, 2: 1})
    assert out.tolist() == [1, -1, -1, 1, -1]
    g = load_golden("deep30")                     # 30 % off-tree labels in the fixture
    _, _, clf = golden_frames(g)
    k2i = {int(c): i for i, c in enumerate(g["clusters"])}
    lab = _prior_labels(clf, g["snv_order"], k2i)
    q0 = g["q_y_init"]
    uniform = np.isclose(q0, 1 / len(k2i)).all(axis=1)
    assert np.array_equal(lab < 0, uniform)
    assert np.array_equal(lab[~uniform], q0[~uniform].argmax(1))


def _toy_fit(lazy):
    from arborist_b200.treefit import TreeFit
    qz = np.array([[0.2, 0.8], [0.5, 0.5], [0.9, 0.1]])
    qy = np.array([[1.0], [1.0]])
    args = ([(0, 1)], 3, -12.5)
    maps = ({"c1": 0, "c2": 1, "c3": 2}, {"s1": 0, "s2": 1}, {0: 0, 1: 1}, {1: 0})
    if lazy:
        calls = []
        return TreeFit.lazy(*args, lambda idx: calls.append(idx) or (qz, qy), *maps), calls
    return TreeFit(*args, qz, qy, *maps), None


def test_treefit_tables_and_map_argmax_first_max_wins():
    f, _ = _toy_fit(False)
    z = f.map_assign_z()
    assert list(z.columns) == ["id", "assignment"]
    assert z["id"].tolist() == ["c1", "c2", "c3"] and z["assignment"].tolist() == [1, 0, 0]
    assert list(f.q_z_df().columns) == ["id", "clone_0", "clone_1"]
    assert list(f.q_y_df().columns) == ["id", "cluster_1"]
    assert "ELBO: -12.50" in str(f) and " 0->1" in str(f)
    assert list(f.cell_entropy().columns) == ["id", "entropy"]


def test_treefit_lazy_loads_once_and_pickles_as_plain_arrays(tmp_path):
    f, calls = _toy_fit(True)
    assert calls == []
    assert f.elbo == -12.5 and calls == []
    assert f.q_z.shape == (3, 2) and calls == [3]
    f.q_y
    assert calls == [3]
    g = pickle.loads(pickle.dumps({3: f}))[3]
    assert type(g).__module__ == "arborist.treefit"          # the path reference pickles use
    assert np.array_equal(g.q_z, f.q_z) and g.map_assign_z()["assignment"].tolist() == [1, 0, 0]
    f.save_tree(tmp_path / "best.txt")
    assert (tmp_path / "best.txt").read_text() == "1 #edges tree 0\n0 1\n"


def test_drop_in_package_names():
    import arborist
    import arborist_b200
    assert arborist.arborist is arborist_b200.arborist
    assert arborist.read_trees is arborist_b200.read_trees
    from arborist.treefit import TreeFit
    assert TreeFit is arborist_b200.TreeFit
    from arborist.arborist_base import rank_trees, find_cell_assignments, build_genotypes  # noqa: F401
    from arborist_b200 import arborist_base
    assert rank_trees is arborist_base.rank_trees


def test_cli_flags_match_reference():
    from arborist_b200.arborist import parse_arguments
    a = parse_arguments(["-R", "r.csv", "-Y", "y.csv", "-T", "t.txt"])
    assert (a.alpha, a.max_iter, a.prior, a.threads, a.edge_delim) == (0.001, 25, 0.7, 10, " ")
    assert not a.add_normal and not a.verbose
    a = parse_arguments(["-R", "r", "-Y", "y", "-T", "t", "--add-normal", "--ranking", "o.csv", "-j", "4",
                         "--q_z", "z.csv", "--q_y", "y.csv", "--pickle", "p.pkl", "-d", "x.png", "-t", "b.txt",
                         "--cell-assign", "c.csv", "--snv-assign", "s.csv", "-v", "--prior", "0.9"])
    assert a.add_normal and a.ranking == "o.csv" and a.threads == 4 and a.prior == 0.9


def test_synth_generator_is_seed_stable():
    from arborist_b200 import synth
    a, b = synth.generate(synth.CONFIGS["tiny"]), synth.generate(synth.CONFIGS["tiny"])
    assert np.array_equal(a.alt, b.alt) and a.trees == b.trees
    g = load_golden("tiny")
    assert np.array_equal(a.snv, g["rc_snv"]) and np.array_equal(a.alt, g["rc_alt"])


def test_read_read_counts_formats(tmp_path):
    """the read-count table as CSV (the reference's format, A.py:1135), Parquet,
    Feather or npz gives the same DataFrame."""
    import pandas as pd
    from arborist_b200.utils import read_read_counts
    from arborist_b200 import synth
    rc = synth.generate(synth.CONFIGS["tiny"]).read_counts_frame()
    rc.to_csv(tmp_path / "rc.csv", index=False)
    rc.to_parquet(tmp_path / "rc.parquet")
    rc.to_feather(tmp_path / "rc.feather")
    np.savez(tmp_path / "rc.npz", **{c: rc[c].to_numpy() for c in rc.columns})
    for name in ("rc.csv", "rc.parquet", "rc.feather", "rc.npz"):
        got = read_read_counts(str(tmp_path / name))
        pd.testing.assert_frame_equal(got[list(rc.columns)], rc, check_dtype=True)
    np.savez(tmp_path / "bad.npz", snv=rc["snv"].to_numpy())
    with pytest.raises((ValueError, KeyError)):
        read_read_counts(str(tmp_path / "bad.npz"))


def test_prepare_tree_shard_matches_full_preparation():
    """The shard-proportional path (``--add-normal`` folded into the parent vectors, no per-tree edge lists) gives
    what add_normal_clone + prepare_trees give on the whole set, for any subset of tree indices; irregular sets,
    forests with two roots and non-integer labels take the general path with the same result."""
    from arborist_b200 import synth
    from arborist_b200.arborist import add_normal_clone, prepare_tree_shard, prepare_trees
    d = synth.generate(synth.CONFIGS["small"])
    full = prepare_trees(add_normal_clone(d.trees))
    T = len(d.trees)
    for idx in (np.arange(T), np.arange(1, T, 3), np.array([T - 1, 0])):
        got = prepare_tree_shard(d.trees, idx, add_normal=True)
        assert got[0] == full[0] and got[1] == full[1] and got[4] == full[4] and got[3] is None
        assert np.array_equal(got[2], full[2][idx])
    with_n = add_normal_clone(d.trees)
    got = prepare_tree_shard(with_n, np.arange(2, T), add_normal=False)
    assert np.array_equal(got[2], full[2][2:])
    # an empty shard (more ranks than trees) still yields the label maps
    got = prepare_tree_shard(d.trees, np.array([], dtype=np.int64), add_normal=True)
    assert got[0] == full[0] and got[2].shape == (0, len(full[0]))
    # ragged / string-labelled sets
    ragged = [[(1, 2), (2, 3)], [(1, 3), (3, 2), (1, 3)]]
    a = prepare_tree_shard(ragged, np.arange(2), add_normal=True)
    b = prepare_trees(add_normal_clone(ragged))
    assert a[0] == b[0] and np.array_equal(a[2], b[2])
    names = [[("a", "b"), ("b", "c")], [("a", "c"), ("c", "b")]]
    a = prepare_tree_shard(names, np.arange(2), add_normal=False)
    assert np.array_equal(a[2], prepare_trees(names)[2])
    with pytest.raises(ValueError, match="same set of clones"):
        prepare_tree_shard([[(1, 2), (2, 3)], [(1, 2), (2, 4)]], np.arange(2), add_normal=True)
    # the mismatch is only seen by the process whose shard holds the bad tree
    prepare_tree_shard([[(1, 2), (2, 3)], [(1, 2), (2, 4)]], np.array([0]), add_normal=True)


def test_lazy_fits_and_streaming_pickle(tmp_path):
    """``all_tree_fits`` behaves like the reference's eager dict; ``to_pickle`` writes the file ``pd.to_pickle`` of
    the eager dict would (plain dict of TreeFits, index dicts shared between fits) without keeping the posteriors
    of more than a few trees alive (A.py:1187-1188 at 1k-10k fits)."""
    import gc
    import os
    import weakref

    from arborist_b200.results import LazyFits, write_fits_pickle
    from arborist_b200.treefit import TreeFit
    n, m, K, T = 40, 60, 5, 2100
    cell = {f"c{i}": i for i in range(n)}
    snv = {i * 3: i for i in range(m)}
    c2i, k2i = {i - 1: i for i in range(K)}, {i: i for i in range(K - 1)}
    alive = []

    def loader(i):
        r = np.random.default_rng(i)
        qz, qy = r.random((n, K)), r.random((m, K - 1))
        alive.append(weakref.ref(qz))
        return qz, qy

    def make(i):
        return TreeFit.lazy(None, i, -float(i), loader, cell, snv, c2i, k2i, tree_loader=lambda j: [(j, 1), (1, 2)])

    fits = LazyFits(T, make)
    assert len(fits) == T and 5 in fits and T not in fits and "x" not in fits
    assert list(fits)[:3] == [0, 1, 2] and list(fits.keys())[-1] == T - 1 and fits.get(T) is None
    with pytest.raises(KeyError):
        fits[T]
    assert fits[7] is fits[7] and fits[7].tree == [(7, 1), (1, 2)]
    path = tmp_path / "fits.pkl"
    fits.to_pickle(path)
    gc.collect()
    assert sum(r() is not None for r in alive) <= 3, "the streaming writer kept posteriors alive"
    back = pd.read_pickle(path)
    assert type(back) is dict and list(back) == list(range(T))
    for i in (0, 1, 999, 1000, 1001, T - 1):
        r = np.random.default_rng(i)
        f = back[i]
        assert np.array_equal(f.q_z, r.random((n, K))) and np.array_equal(f.q_y, r.random((m, K - 1)))
        assert f.elbo == -float(i) and f.tree == [(i, 1), (1, 2)] and f.tree_idx == i
        assert f.cell_to_idx is back[0].cell_to_idx and f.snv_to_idx is back[0].snv_to_idx
        assert list(f.__dict__)[:9] == ["tree", "tree_idx", "elbo", "q_z", "q_y", "cell_to_idx", "snv_to_idx",
                                        "clone_to_idx", "cluster_to_idx"]
    eager = {i: make(i) for i in range(T)}
    pd.to_pickle(eager, tmp_path / "eager.pkl")
    assert os.path.getsize(path) < 1.05 * os.path.getsize(tmp_path / "eager.pkl")
    plain = pickle.loads(pickle.dumps(LazyFits(6, make)))          # pd.to_pickle(all_fits) still works
    assert type(plain) is dict and len(plain) == 6 and plain[3].elbo == -3.0
    write_fits_pickle(LazyFits(4, make), tmp_path / "fits.pkl.gz")
    assert len(pd.read_pickle(tmp_path / "fits.pkl.gz")) == 4


def test_pyproject_console_script():
    """The reference installs ``arborist = "arborist.arborist:main"`` (pyproject.toml:34-35)."""
    import importlib
    import os
    import tomllib

    from conftest import ROOT
    cfg = tomllib.load(open(os.path.join(ROOT, "pyproject.toml"), "rb"))
    target = cfg["project"]["scripts"]["arborist"]
    assert target == "arborist.arborist:main"
    mod, fn = target.split(":")
    assert callable(getattr(importlib.import_module(mod), fn))


def test_pick_devices_counts_lists_and_bad_ordinals(monkeypatch):
    """``devices`` / ``ARBORIST_DEVICES``: "all", a count (int or bare integer string), or a list of ordinals; an
    ordinal the process cannot see is a ValueError here, not a CUDA error later."""
    import importlib
    from arborist_b200 import engine
    A = importlib.import_module("arborist_b200.arborist")
    monkeypatch.setattr(engine, "device_count", lambda: 4)
    monkeypatch.delenv("ARBORIST_DEVICES", raising=False)
    pick = A._pick_devices
    assert pick(None, "all", 1000) == [0, 1, 2, 3]
    assert pick(None, 2, 1000) == [0, 1] and pick(None, "2", 1000) == [0, 1]      # the string means the same count
    assert pick(None, "1", 1000) == [0]
    assert pick(None, "3,", 1000) == [3] and pick(None, "0, 2", 1000) == [0, 2] and pick(None, [1, 3], 1000) == [1, 3]
    assert pick(None, 9, 1000) == [0, 1, 2, 3]                                     # a count is capped
    assert pick(None, "all", 2) == [0, 1]                                          # never more GPUs than trees
    assert pick(2, "all", 1000) == [2]                                             # an explicit device wins
    monkeypatch.setenv("ARBORIST_DEVICES", "1")
    assert pick(None, None, 1000) == [0]
    monkeypatch.setenv("ARBORIST_DEVICES", "1,2")
    assert pick(None, None, 1000) == [1, 2]
    for bad in ("4,", [0, 7], "-1,"):
        with pytest.raises(ValueError, match="CUDA device"):
            pick(None, bad, 1000)
    with pytest.raises(ValueError, match="empty"):
        pick(None, ",", 1000)


def test_seam_fingerprint_sees_every_element():
    """The per-tree seam keeps the device matrix between calls and recognises its inputs by content: an in-place edit
    of ONE element anywhere (the old strided sample missed most positions) must change the fingerprint."""
    from arborist_b200.arborist import _fingerprint
    rng = np.random.default_rng(0)
    for a in (rng.random(100_003), rng.integers(0, 1000, 70_001).astype(np.int32), rng.random((3001, 7)),
              np.arange(5, dtype=np.int8), np.zeros(0)):
        base = _fingerprint(a)
        assert _fingerprint(a) == base and _fingerprint(a.copy())[1:] == base[1:]
        flat = a.reshape(-1)
        for pos in ([0, 1, flat.shape[0] // 2 + 1, flat.shape[0] - 2, flat.shape[0] - 1] if flat.shape[0] else []):
            old = flat[pos]
            flat[pos] = old + 1
            assert _fingerprint(a) != base, pos
            flat[pos] = old
        assert _fingerprint(a) == base
    b = rng.random(4096 * 5 + 17)
    base = _fingerprint(b)
    for i, j in ((1, 2), (3, 4096 * 2 + 3), (10, 4096 * 4 + 11), (4096 * 5 + 1, 4096 * 5 + 9)):
        b[i], b[j] = b[j], b[i]                   # an exchange keeps every plain sum / xor of the values
        assert _fingerprint(b) != base, (i, j)
        b[i], b[j] = b[j], b[i]
    assert _fingerprint(b) == base
