"""Pins ``oracle/`` (C restatement + numpy host steps) against the golden vectors that
``oracle/gen_golden.py`` recorded from the UNMODIFIED reference (kernel, driver, ingest
and API level).  CPU only."""
import numpy as np
import pytest

from conftest import golden_frames, load_golden, synth_frames
from oracle import oracle as orc

SMALL = ["tiny", "small", "small_nonormal", "deep30", "k2"]


def _prep(g):
    trees, rc, cl = golden_frames(g)
    return orc.prepare(trees, rc, cl, alpha=float(g["alpha"]), gamma=float(g["gamma"]),
                       add_normal=bool(g["add_normal"]))


@pytest.mark.parametrize("name", SMALL)
def test_ingest_matches_reference(name):
    g = load_golden(name)
    p = _prep(g)
    assert p.n_cells == int(g["n_cells"]) and p.n_snvs == int(g["n_snvs"])
    assert p.n_clones == int(g["n_clones"])
    assert np.array_equal(np.array(list(p.cell_to_idx)), g["cell_order"])
    assert np.array_equal(np.array(list(p.snv_to_idx)), g["snv_order"])
    assert np.array_equal(np.array(list(p.clone_to_idx)), g["clones"])
    assert np.array_equal(np.array(list(p.cluster_to_idx)), g["clusters"])
    assert np.array_equal(p.cell_ptr, g["cell_ptr"]) and np.array_equal(p.snv_ptr, g["snv_ptr"])
    assert np.array_equal(p.q_y_init, g["q_y_init"])
    # within-row order is unspecified (unstable argsort, A.py:448): compare as per-row sets
    for ptr, idx, ll, gidx, gll in ((p.cell_ptr, p.snv_idx_cs, p.ll_cs, g["snv_idx_cs"], g["ll_cs"]),
                                    (p.snv_ptr, p.cell_idx_ss, p.ll_ss, g["cell_idx_ss"], g["ll_ss"])):
        a = np.stack([np.repeat(np.arange(len(ptr) - 1), np.diff(ptr)), idx, ll[0::2], ll[1::2]], 1)
        b = np.stack([np.repeat(np.arange(len(ptr) - 1), np.diff(ptr)), gidx, gll[0::2], gll[1::2]], 1)
        assert np.array_equal(a[np.lexsort(a.T[::-1])], b[np.lexsort(b.T[::-1])])


@pytest.mark.parametrize("name", SMALL)
def test_presence_matches_reference(name):
    g = load_golden(name)
    p = _prep(g)
    for t, tree in enumerate(p.tree_list):
        assert np.array_equal(orc.enumerate_presence(tree, p.clone_to_idx, p.cluster_to_idx),
                              g["presence"][t])


@pytest.mark.parametrize("name", SMALL)
def test_kernels_match_reference(name):
    g = load_golden(name)
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    P = g["presence"][0]
    lqy = orc.log_array(g["k_qy_in"])
    assert np.isneginf(lqy[0, 1:]).all()
    qz = orc.compute_q_z_sparse(g["cell_ptr"], g["snv_idx_cs"], g["ll_cs"], lqy, P, n, K)
    np.testing.assert_allclose(qz, g["k_qz"], rtol=1e-12, atol=1e-300)
    lqz = orc.log_array(g["k_qz"])
    qy = orc.compute_q_y_sparse(g["snv_ptr"], g["cell_idx_ss"], g["ll_ss"], lqz, P, m, K,
                                orc.log_array(g["q_y_init"]))
    np.testing.assert_allclose(qy, g["k_qy"], rtol=1e-12, atol=1e-300)
    lik = orc.compute_likelihood_sparse(g["cell_ptr"], g["snv_idx_cs"], g["ll_cs"], lqy, lqz, P, K, n)
    assert lik == pytest.approx(float(g["k_lik"]), rel=1e-12)
    assert orc.compute_entropy(g["k_qz"]) == pytest.approx(float(g["k_ent"]), rel=1e-13)
    assert orc.compute_kl_divergence(g["k_qy"], g["q_y_init"]) == pytest.approx(float(g["k_kl"]),
                                                                              rel=1e-13)


@pytest.mark.parametrize("name", SMALL)
def test_driver_matches_reference(name):
    """ELBO, iteration count, trace and best posteriors of run_variational_inference."""
    g = load_golden(name)
    p = _prep(g)
    for t, tree in enumerate(p.tree_list):
        elbo, qz, qy, it, trace = orc.fit_tree(p, tree, int(g["max_iter"]), float(g["tol"]),
                                               return_trace=True)
        assert it == int(g["iters"][t])
        np.testing.assert_allclose(trace, g["traces"][t, : it + 1], rtol=1e-11)
        assert elbo == pytest.approx(float(g["elbo"][t]), rel=1e-11)
        np.testing.assert_allclose(qz, g["q_z"][t], rtol=1e-9, atol=1e-300)
        np.testing.assert_allclose(qy, g["q_y"][t], rtol=1e-9, atol=1e-300)
        assert np.array_equal(qz.argmax(1), g["map_z"][t])
        assert np.array_equal(qy.argmax(1), g["map_y"][t])


@pytest.mark.parametrize("name", SMALL)
def test_hard_assignment_mode_matches_reference(name):
    g = load_golden(name)
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    lik, qz, qy = orc.run_simple_max_likelihood(
        g["presence"][0], g["ll_cs"], g["ll_ss"], g["cell_idx_ss"], g["snv_idx_cs"], n, m, K,
        g["q_y_init"], g["cell_ptr"], g["snv_ptr"])
    assert lik == pytest.approx(float(g["ml_lik"]), rel=1e-12)
    np.testing.assert_allclose(qz, g["ml_qz"], rtol=1e-12, atol=1e-300)
    assert np.array_equal(qy, g["ml_qy"])


def test_api_level_example_standin_subset():
    """BASELINE config 1 stand-in: ranking of a tree subset + best-tree MAP labels."""
    g = load_golden("example_standin")
    trees, rc, cl = golden_frames(g)
    sub = [0, 7, 19]
    lik, best, fits, prob = orc.arborist_oracle(trees, rc, cl, alpha=float(g["alpha"]),
                                                max_iter=int(g["max_iter"]), tolerance=float(g["tol"]),
                                                gamma=float(g["gamma"]),
                                                add_normal=bool(g["add_normal"]), trees=set(sub))
    for t in sub:
        assert lik[t] == pytest.approx(float(g["elbo"][t]), rel=1e-11)
        assert fits[t][3] == int(g["iters"][t])
        assert np.array_equal(fits[t][1].argmax(1), g["map_z"][t])
        assert np.array_equal(fits[t][2].argmax(1), g["map_y"][t])
    assert best == 0 == int(g["best_idx"])
    np.testing.assert_allclose(fits[0][1], g["q_z_best"], rtol=1e-9, atol=1e-300)


def test_config2_subset_from_seeded_generator():
    """The synthetic generator reproduces the inputs the golden run used (seed 1002)."""
    g = load_golden("config2")
    trees, rc, cl = synth_frames("config2")
    assert np.array_equal(np.array(trees), g["trees"])
    sub = [0, 24]
    lik, _, fits, prob = orc.arborist_oracle(trees, rc, cl, max_iter=25, tolerance=1.0, gamma=0.7,
                                             add_normal=True, trees=set(sub))
    assert prob.n_snvs == int(g["n_snvs"]) and prob.ll_cs.shape[0] == 2 * int(g["nnz"])
    for t in sub:
        assert lik[t] == pytest.approx(float(g["elbo"][t]), rel=1e-11)
        assert fits[t][3] == int(g["iters"][t])
        assert np.array_equal(fits[t][1].argmax(1), g["map_z"][t])


def test_config5_full_shape_one_tree():
    """BASELINE config 5 at its real shape (5k x 10k, K = 30 deep trees, 30 % off-tree SNVs -> uniform prior rows):
    the oracle against the live reference's run of the same tree (tests/golden/config5_full.npz)."""
    g = load_golden("config5_full")
    sub_all = [int(i) for i in g["tree_subset"]]
    trees, rc, cl = synth_frames("config5", tree_subset=sub_all)
    assert np.array_equal(np.array(trees), g["trees"])
    k = int(np.argmin(g["iters"]))                        # the shortest fit keeps the CPU suite short
    lik, _, fits, prob = orc.arborist_oracle(trees, rc, cl, max_iter=25, tolerance=1.0, gamma=0.7,
                                             add_normal=True, trees={k})
    assert (prob.n_cells, prob.n_snvs, prob.n_clones) == (int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"]))
    assert lik[k] == pytest.approx(float(g["elbo"][k]), rel=1e-11)
    assert fits[k][3] == int(g["iters"][k])
    assert np.array_equal(fits[k][1].argmax(1), g["map_z"][k])
    assert np.array_equal(fits[k][2].argmax(1), g["map_y"][k])
    r = g["q_z_rows"].shape[1]
    np.testing.assert_allclose(fits[k][1][:r], g["q_z_rows"][k], rtol=1e-9, atol=1e-300)
