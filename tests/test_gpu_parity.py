"""GPU parity tests: the CUDA path (through the C ABI) against the golden vectors recorded
from the unmodified reference and against the CPU oracle on fresh seeded inputs.

Tolerances (BASELINE.json north_star): tree ranking, iteration counts and MAP assignments
bit-exact; ELBOs and posteriors within 1e-6 relative in fp64 -- the tests ask for 1e-9
(observed ~1e-13) so a regression shows long before the contract is at risk."""
import os

import numpy as np
import pandas as pd
import pytest

from conftest import golden_frames, load_golden, synth_frames

pytestmark = pytest.mark.gpu

ELBO_RTOL = 1e-9
# posteriors: observed agreement with the reference is ~1e-13 relative on entries above 1e-6 and a few 1e-16
# absolute below; the tests ask for 100x that (north_star allows 1e-6)
Q_RTOL, Q_ATOL = 1e-10, 1e-13
MAP_TIE_ROWS = {}                  # test id -> rows that took the tie branch of assert_map_equal


def assert_map_equal(q, q_ref, max_ties=2, where="?"):
    """MAP labels must be identical wherever the reference's own maximum is decided by more
    than rounding: rows whose top two reference posteriors differ by < 1e-9 relative are exact
    mathematical ties (e.g. sibling leaves with identical evidence) that the reference breaks
    by the last bit of its log-sum-exp; there the chosen label must still be one of the tied.
    Such rows are COUNTED and bounded (``max_ties`` per matrix), so a real regression cannot hide here."""
    a, b = q.argmax(1), q_ref.argmax(1)
    diff = np.nonzero(a != b)[0]
    for i in diff:
        top = q_ref[i, b[i]]
        assert abs(q_ref[i, a[i]] - top) <= 1e-9 * top, f"MAP differs on row {i} beyond a tie"
    MAP_TIE_ROWS[where] = MAP_TIE_ROWS.get(where, 0) + len(diff)
    assert len(diff) <= max_ties, f"{len(diff)} MAP rows differ from the reference (ties only, but too many): {where}"


@pytest.fixture(scope="module")
def eng_mod():
    from arborist_b200 import engine
    assert engine.device_count() > 0, "no CUDA device: the gpu tests must run on the B200 box"
    return engine


def _row_sets(ptr, idx, ll):
    a = np.stack([np.repeat(np.arange(len(ptr) - 1), np.diff(ptr)), idx, ll[0::2], ll[1::2]], 1)
    return a[np.lexsort(a.T[::-1])]


def _coo_engine(eng_mod, g, with_scipy=True):
    from arborist_b200.arborist import _install_scipy_table
    trees, rc, cl = golden_frames(g)
    snv_codes, snv_u = pd.factorize(rc["snv"].to_numpy())
    cell_codes, cell_u = pd.factorize(rc["cell"].to_numpy())
    e = eng_mod.Engine(0)
    e.ingest_coo(cell_codes, snv_codes, rc["alt"].to_numpy(), rc["total"].to_numpy(), len(cell_u),
                 len(snv_u), float(g["alpha"]))
    if with_scipy:
        _install_scipy_table(e, float(g["alpha"]))
    return e, cell_u, snv_u


SMALL = ["tiny", "small", "small_nonormal", "deep30", "k2"]


@pytest.mark.parametrize("name", SMALL)
def test_ingest_matches_reference(eng_mod, name):
    """a1-a5: filter, index order, CSR/CSC content, log-likelihood table, prior."""
    g = load_golden(name)
    e, cell_u, snv_u = _coo_engine(eng_mod, g, with_scipy=False)
    n, m, nnz, n_pairs = e.dims()
    assert (n, m, nnz) == (int(g["n_cells"]), int(g["n_snvs"]), int(g["nnz"]))
    co, so = e.orders()
    assert np.array_equal(np.asarray(cell_u)[co], g["cell_order"])
    assert np.array_equal(np.asarray(snv_u)[so], g["snv_order"])
    # device lgamma table vs scipy: close; after installing scipy's doubles: bit-equal
    from scipy.stats import binom
    alt, tot = e.count_pairs()
    a = float(g["alpha"])
    lp, la = e.loglik_table()
    np.testing.assert_allclose(lp, binom.logpmf(alt, tot, 0.5 - a + 0.5 * a / 3), rtol=1e-13, atol=1e-15)
    np.testing.assert_allclose(la, binom.logpmf(alt, tot, a / 3), rtol=1e-13, atol=1e-15)
    from arborist_b200.arborist import _install_scipy_table
    _install_scipy_table(e, a)
    ptr, idx, ll = e.export_sparse(False)
    assert np.array_equal(ptr, g["cell_ptr"])
    assert np.array_equal(_row_sets(ptr, idx, ll), _row_sets(g["cell_ptr"], g["snv_idx_cs"], g["ll_cs"]))
    ptr, idx, ll = e.export_sparse(True)
    assert np.array_equal(ptr, g["snv_ptr"])
    assert np.array_equal(_row_sets(ptr, idx, ll), _row_sets(g["snv_ptr"], g["cell_idx_ss"], g["ll_ss"]))
    # prior from labels
    from arborist_b200.arborist import _prior_labels
    _, _, cl = golden_frames(g)
    k2i = {int(c): i for i, c in enumerate(g["clusters"])}
    lab = _prior_labels(cl, g["snv_order"], k2i)
    e.set_prior_labels(lab, len(k2i), float(g["gamma"]))
    assert np.array_equal(e.prior(len(k2i)), g["q_y_init"])
    e.close()


@pytest.mark.parametrize("name", SMALL)
def test_kernel_hooks_match_reference(eng_mod, name):
    """a8-a11 on explicit arrays, matrix ingested through the seam-level CSR entry point."""
    g = load_golden(name)
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    e = eng_mod.Engine(0)
    e.ingest_csr(g["cell_ptr"], g["snv_idx_cs"], g["ll_cs"], n, m)
    e.set_prior(g["q_y_init"])
    P = g["presence"][0]
    qz = e.qz_update(P, g["k_qy_in"])
    np.testing.assert_allclose(qz, g["k_qz"], rtol=Q_RTOL, atol=Q_ATOL)
    assert np.array_equal(qz.argmax(1), g["k_qz"].argmax(1))
    qy = e.qy_update(P, g["k_qz"])
    np.testing.assert_allclose(qy, g["k_qy"], rtol=Q_RTOL, atol=Q_ATOL)
    assert np.array_equal(qy.argmax(1), g["k_qy"].argmax(1))
    lik, ent, kl = e.elbo_terms(P, g["k_qz"], g["k_qy_in"])
    assert lik == pytest.approx(float(g["k_lik"]), rel=1e-11)
    assert ent == pytest.approx(float(g["k_ent"]), rel=1e-12)
    _, _, kl2 = e.elbo_terms(P, g["k_qz"], g["k_qy"])
    assert kl2 == pytest.approx(float(g["k_kl"]), rel=1e-12)
    e.close()


def _assert_golden_map(q, ref_idx, max_tie_rows, where):
    """MAP against the reference's own argmax.  Identical, except for at most ``max_tie_rows`` rows in which the
    reference's label and ours carry the SAME posterior to 1e-9 relative (exact mathematical ties, e.g. a cell whose
    reads cannot tell sibling clones apart: q = 1/3, 1/3, 1/3 to 13 digits) -- the reference breaks those by the
    last bit of its log-sum-exp.  Only config 4 (100,000 cells after 2 iterations) is given such an allowance."""
    a = q.argmax(1)
    diff = np.nonzero(a != ref_idx)[0]
    if len(diff) == 0:
        return
    assert len(diff) <= max_tie_rows, f"{where}: {len(diff)} MAP rows differ from the reference"
    for i in diff:
        assert abs(q[i, a[i]] - q[i, ref_idx[i]]) <= 1e-9 * q[i, a[i]], f"{where}: MAP differs on row {i} beyond a tie"
    MAP_TIE_ROWS["golden_" + where] = MAP_TIE_ROWS.get("golden_" + where, 0) + len(diff)


def _check_api(g, lik, best, fits, full_q, max_tie_rows=0):
    T = len(g["elbo"])
    elbo = np.array([lik[t] for t in range(T)])
    np.testing.assert_allclose(elbo, g["elbo"], rtol=ELBO_RTOL)
    # ranking identical (ties keep index order on both sides)
    assert np.array_equal(np.argsort(-elbo, kind="stable"), np.argsort(-g["elbo"], kind="stable"))
    assert best.tree_idx == int(g["best_idx"])
    for t in range(T):
        f = fits[t]
        z = f.map_assign_z()["assignment"].to_numpy()
        y = f.map_assign_y()["assignment"].to_numpy()
        if max_tie_rows == 0:
            assert np.array_equal(z, g["clones"][g["map_z"][t]]), f"MAP z differs for tree {t}"
            assert np.array_equal(y, g["clusters"][g["map_y"][t]]), f"MAP y differs for tree {t}"
        else:
            if not np.array_equal(z, g["clones"][g["map_z"][t]]):
                _assert_golden_map(f.q_z, g["map_z"][t], max_tie_rows, f"{str(g['name'])} tree {t} z")
            if not np.array_equal(y, g["clusters"][g["map_y"][t]]):
                _assert_golden_map(f.q_y, g["map_y"][t], max_tie_rows, f"{str(g['name'])} tree {t} y")
    if full_q:
        for t in range(T):
            np.testing.assert_allclose(fits[t].q_z, g["q_z"][t], rtol=Q_RTOL, atol=Q_ATOL)
            np.testing.assert_allclose(fits[t].q_y, g["q_y"][t], rtol=Q_RTOL, atol=Q_ATOL)
    else:
        r = g["q_z_rows"].shape[1]
        for t in range(0, T, max(1, T // 8)):
            np.testing.assert_allclose(fits[t].q_z[:r], g["q_z_rows"][t], rtol=Q_RTOL, atol=Q_ATOL)
            np.testing.assert_allclose(fits[t].q_y[:r], g["q_y_rows"][t], rtol=Q_RTOL, atol=Q_ATOL)
        if "q_z_best" in g.files:                  # config 4 leaves the winner's 24 MB of posteriors out of the fixture
            np.testing.assert_allclose(best.q_z, g["q_z_best"], rtol=Q_RTOL, atol=Q_ATOL)
            np.testing.assert_allclose(best.q_y, g["q_y_best"], rtol=Q_RTOL, atol=Q_ATOL)


def _run_api(g, trees, rc, cl, **kw):
    from arborist_b200 import arborist
    from arborist_b200.arborist import arborist as fn
    lik, best, fits = arborist(trees, rc, cl, alpha=float(g["alpha"]), max_iter=int(g["max_iter"]),
                               tolerance=float(g["tol"]), gamma=float(g["gamma"]),
                               add_normal=bool(g["add_normal"]), **kw)
    iters = fn.last_stats["iters"]
    assert np.array_equal(iters, g["iters"]), "iteration counts differ from the reference"
    assert list(best.cell_to_idx) == list(g["cell_order"])
    assert list(best.snv_to_idx) == list(g["snv_order"])
    assert list(best.clone_to_idx) == list(g["clones"])
    assert list(best.cluster_to_idx) == list(g["clusters"])
    return lik, best, fits


@pytest.mark.parametrize("name", SMALL)
def test_api_matches_reference_small(eng_mod, name):
    g = load_golden(name)
    trees, rc, cl = golden_frames(g)
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=True)


def test_api_example_standin(eng_mod):
    """BASELINE config 1 (bundled trees + clustering, synthetic counts; no --add-normal)."""
    g = load_golden("example_standin")
    trees, rc, cl = golden_frames(g)
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=False)


def test_api_config2(eng_mod):
    """BASELINE config 2: 1k cells x 5k SNVs, 8 clones, 100 trees."""
    g = load_golden("config2")
    trees, rc, cl = synth_frames("config2")
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=False)


def test_api_config3_head(eng_mod):
    """BASELINE config 3 shape (10k x 20k, 12 clones): the 4 trees the reference was run on."""
    g = load_golden("config3_head")
    trees, rc, cl = synth_frames("config3", tree_subset=[int(i) for i in g["tree_subset"]])
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=False)


def _check_traces(g, trees, eng):
    """Driver level on the device matrix the API call left behind: every iteration's ELBO (A.py:801-803)."""
    from arborist_b200.arborist import prepare_tree_shard
    c2i, _, parents, _, normal = prepare_tree_shard(trees, np.arange(len(trees)), add_normal=True)
    elbo, iters, tr = eng.fit(parents=parents, n_clones=len(c2i), normal_idx=normal, max_iter=int(g["max_iter"]),
                              tolerance=float(g["tol"]), keep_posteriors="none", trace=True)
    assert np.array_equal(iters, g["iters"])
    for t in range(len(trees)):
        assert np.isnan(tr[t, iters[t] + 1:]).all()
    return elbo, tr


def test_api_config3_converged(eng_mod):
    """BASELINE config 3 shape, the first 8 candidate trees fitted TO CONVERGENCE (7 to 25 iterations, three of them
    stopped by max_iter) -- the late-iteration regime of the loop (A.py:767-814: peaked posteriors, snapshot on
    strict improvement, convergence against the best ELBO).  Fixture generated by the C oracle, which the
    live-reference goldens pin (oracle/gen_oracle_golden.py; SURVEY.md 8(d) names these 8 trees): ELBOs and every
    iteration's ELBO to 1e-9, iteration counts, ranking and MAP labels exact, 32 recorded rows of each posterior."""
    from arborist_b200.arborist import arborist as fn
    g = load_golden("config3_conv")
    assert str(g["source"]) == "oracle" and int(g["iters"].max()) == 25 and int(g["iters"].min()) < 10
    trees, rc, cl = synth_frames("config3", tree_subset=[int(i) for i in g["tree_subset"]])
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=False)
    elbo, tr = _check_traces(g, trees, fn.last_stats["engine"]())
    np.testing.assert_allclose(tr, g["traces"], rtol=ELBO_RTOL)          # NaN padding included
    assert np.array_equal(elbo, np.array([lik[t] for t in range(len(trees))]))


def test_api_config5_full_shape(eng_mod):
    """BASELINE config 5 at its REAL shape: 5,000 cells x 10,000 SNVs, K = 30 deep trees, 30 % of the SNVs carry
    an off-tree label (uniform prior rows); the 7 first trees + the generating tree, as the unmodified reference
    fitted them (oracle/gen_golden.py config5_full): ELBOs 1e-9, iteration counts, ranking and MAP exact, the
    recorded posterior rows and the winner's full posteriors."""
    g = load_golden("config5_full")
    sub = [int(i) for i in g["tree_subset"]]
    trees, rc, cl = synth_frames("config5", tree_subset=sub)
    lik, best, fits = _run_api(g, trees, rc, cl)
    _check_api(g, lik, best, fits, full_q=False)
    assert sub[best.tree_idx] == int(g["true_tree_index"])           # the generating tree wins, as in the reference


def test_api_config4_head(eng_mod):
    """BASELINE config 4 at its REAL shape: 100,000 cells x 50,000 SNVs, nnz 99.0 M, K = 20 -- the shape with
    28 / 56 panels, 782 / 391 tiles per warp and split sweep launches.  The first 2 candidate trees with max_iter
    capped at 2, as the unmodified reference fitted them in the build container (36 min of numba on 8 threads,
    oracle/gen_golden.py config4_head; SURVEY.md 8(d) names these 2 trees as the parity sample): ELBOs to 1e-9,
    ELBO traces, iteration counts, ranking, MAP of all 50,000 SNVs and of the 100,000 cells identical except for
    exact ties (54 cells of tree 0 with q = 1/3, 1/3, 1/3 to 13 digits; counted, bounded at 60), 32 recorded rows
    of every posterior.  Generating the workload takes about a minute of host time."""
    from arborist_b200.arborist import arborist as fn
    g = load_golden("config4_head")
    sub = [int(i) for i in g["tree_subset"]]
    trees, rc, cl = synth_frames("config4", tree_subset=sub)
    assert np.array_equal(np.array(trees), g["trees"]) and len(rc) > 9e7
    lik, best, fits = _run_api(g, trees, rc, cl, device=0)
    # 54 of tree 0's 100,000 cells are three-way ties (q = 1/3 each to 1e-13) after 2 iterations: allowance = seen + 10 %
    _check_api(g, lik, best, fits, full_q=False, max_tie_rows=60)
    eng = fn.last_stats["engine"]()
    assert eng.dims() == (int(g["n_cells"]), int(g["n_snvs"]), int(g["nnz"]), eng.dims()[3])
    pf = eng.last_fit_profile()
    assert pf["panels_z"] >= 28 and pf["panels_y"] >= 56
    # driver level on the same device matrix: every iteration's ELBO (A.py:801-803)
    from arborist_b200.arborist import prepare_tree_shard
    c2i, _, parents, _, normal = prepare_tree_shard(trees, np.arange(len(trees)), add_normal=True)
    elbo, iters, tr = eng.fit(parents=parents, n_clones=len(c2i), normal_idx=normal, max_iter=int(g["max_iter"]),
                              tolerance=float(g["tol"]), keep_posteriors="none", trace=True)
    # The reference adds its 3.8e10 likelihood terms one by one into 8 per-thread doubles (A.py:409-419): at this
    # size that summation carries ~1e-9 relative noise of its own while q(y) is still the flat prior: the C oracle
    # reproduces the reference's ELBO_0 bit for bit with 8 threads (-50269807.296489425) and gives -50269807.1699
    # with 3 threads, 2.5e-9 away (profiles/r2_config4_elbo0_threads.md).  The GPU value differs by 2.8e-9; ELBO_1
    # and ELBO_2 by 1e-10, inside ELBO_RTOL.  So the first pass is compared at 1e-8 (north_star: 1e-6).
    np.testing.assert_allclose(tr[:, 1:], g["traces"][:, 1:], rtol=ELBO_RTOL)
    np.testing.assert_allclose(tr[:, 0], g["traces"][:, 0], rtol=1e-8)
    assert np.array_equal(elbo, np.array([lik[t] for t in range(len(trees))]))
    del eng


def test_api_config4_converged(eng_mod):
    """BASELINE config 4 at its REAL shape, fitted TO CONVERGENCE: the generating tree and tree 0 of the candidate
    set (as many of the two as oracle/gen_oracle_golden.py had finished; the fixture says which).  This is the
    late-iteration regime that config4_head (2 iterations of the numba reference) cannot reach at 28 / 56 panels:
    peaked posteriors, snapshot on strict improvement, convergence against the best ELBO (A.py:767-814).  Fixture
    generated by the C oracle with 8 threads -- the thread count at which it reproduces the reference's own ELBO_0
    at this shape bit for bit (profiles/r2_config4_elbo0_threads.md).  ELBOs and every later iteration's ELBO to
    1e-9 (ELBO_0 to 1e-8, see test_api_config4_head), iteration counts and ranking exact, MAP exact up to counted
    exact ties, 32 recorded rows of every posterior."""
    from arborist_b200.arborist import arborist as fn
    g = load_golden("config4_conv")
    assert str(g["source"]) == "oracle" and int(g["omp_threads"]) == 8
    sub = [int(i) for i in g["tree_subset"]]
    assert sub[0] == int(g["true_tree_index"])
    trees, rc, cl = synth_frames("config4", tree_subset=sub)
    assert np.array_equal(np.array(trees), g["trees"])
    lik, best, fits = _run_api(g, trees, rc, cl, device=0)
    _check_api(g, lik, best, fits, full_q=False, max_tie_rows=60)
    eng = fn.last_stats["engine"]()
    assert eng.dims()[:3] == (int(g["n_cells"]), int(g["n_snvs"]), int(g["nnz"]))
    elbo, tr = _check_traces(g, trees, eng)
    np.testing.assert_allclose(tr[:, 1:], g["traces"][:, 1:], rtol=ELBO_RTOL)
    np.testing.assert_allclose(tr[:, 0], g["traces"][:, 0], rtol=1e-8)
    assert np.array_equal(elbo, np.array([lik[t] for t in range(len(trees))]))
    del eng


@pytest.mark.parametrize("name", ["tiny", "deep30"])
def test_elbo_trace_and_presence(eng_mod, name):
    """Driver level: every iteration's ELBO (A.py:801-803) and the device-built genotype matrix."""
    from arborist_b200.arborist import add_normal_clone, prepare_trees
    g = load_golden(name)
    e, _, _ = _coo_engine(eng_mod, g)
    e.set_prior(g["q_y_init"])
    trees, _, _ = golden_frames(g)
    if bool(g["add_normal"]):
        trees = add_normal_clone(trees)
    c2i, k2i, parents, presence, normal = prepare_trees(trees)
    assert presence is None
    K = len(c2i)
    elbo, iters, tr = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=int(g["max_iter"]),
                            tolerance=float(g["tol"]), keep_posteriors="all", trace=True)
    assert np.array_equal(iters, g["iters"])
    for t in range(len(trees)):
        assert np.array_equal(e.get_presence(t), g["presence"][t])
        np.testing.assert_allclose(tr[t, : iters[t] + 1], g["traces"][t, : iters[t] + 1], rtol=ELBO_RTOL)
        assert np.isnan(tr[t, iters[t] + 1:]).all()
        z, y = e.get_map(t)
        assert np.array_equal(z, g["map_z"][t]) and np.array_equal(y, g["map_y"][t])
    # explicit genotype matrices give the same bits as parent vectors
    elbo2, iters2 = e.fit(presence=g["presence"].astype(np.uint8), n_clones=K, max_iter=int(g["max_iter"]),
                          tolerance=float(g["tol"]), keep_posteriors="none")
    assert np.array_equal(elbo, elbo2) and np.array_equal(iters, iters2)
    # posteriors refit on demand == resident posteriors
    qz_refit, qy_refit = e.get_fit(1)
    e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=int(g["max_iter"]),
          tolerance=float(g["tol"]), keep_posteriors="all")
    qz, qy = e.get_fit(1)
    assert np.array_equal(qz, qz_refit) and np.array_equal(qy, qy_refit)
    e.close()


def test_seam_run_variational_inference(eng_mod):
    """The 13-argument seam `run = run_variational_inference` (A.py:1076, A.py:679-816)."""
    from arborist_b200 import run_variational_inference
    g = load_golden("small")
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    for t in (0, 1, 5):
        elbo, qz, qy = run_variational_inference(
            g["presence"][t], g["ll_cs"], g["ll_ss"], g["cell_idx_ss"], g["snv_idx_cs"], n, m, K,
            g["q_y_init"], g["cell_ptr"], g["snv_ptr"], int(g["max_iter"]), float(g["tol"]))
        assert elbo == pytest.approx(float(g["elbo"][t]), rel=ELBO_RTOL)
        np.testing.assert_allclose(qz, g["q_z"][t], rtol=Q_RTOL, atol=Q_ATOL)
        np.testing.assert_allclose(qy, g["q_y"][t], rtol=Q_RTOL, atol=Q_ATOL)


@pytest.mark.parametrize("name", SMALL)
def test_seam_run_simple_max_likelihood(eng_mod, name):
    """Hard-assignment mode behind the same 13-argument seam (A.py:598-676): likelihood under the
    MAP cluster of the prior and the MAP clone of q(z); returns the soft q(z) and the one-hot q(y)."""
    from arborist_b200 import run_simple_max_likelihood
    g = load_golden(name)
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    lik, qz, qy = run_simple_max_likelihood(
        g["presence"][0], g["ll_cs"], g["ll_ss"], g["cell_idx_ss"], g["snv_idx_cs"], n, m, K,
        g["q_y_init"], g["cell_ptr"], g["snv_ptr"])
    assert lik == pytest.approx(float(g["ml_lik"]), rel=ELBO_RTOL)
    np.testing.assert_allclose(qz, g["ml_qz"], rtol=Q_RTOL, atol=Q_ATOL)
    assert np.array_equal(qy, g["ml_qy"])                      # one-hot: bit-exact
    # a one-hot q(y) gives clones whose ancestor sets differ only in clusters without SNVs EXACTLY equal evidence:
    # many exact ties on the deep 30-clone trees (counted, not part of the module total)
    assert_map_equal(qz, g["ml_qz"], max_ties=16, where=f"ml_{name}")


def test_ml_mode_batched_matches_oracle(eng_mod):
    """ARB_MODE_ML over a whole candidate set in one call vs the oracle's per-tree hard scorer."""
    from oracle import oracle as orc
    g = load_golden("small")
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    e, _, _ = _coo_engine(eng_mod, g)
    e.set_prior(g["q_y_init"])
    P = np.ascontiguousarray(g["presence"] != 0, np.uint8)
    lik, iters = e.fit(presence=P, n_clones=K, max_iter=10, tolerance=1.0, mode="ml", keep_posteriors="all")
    assert np.all(iters == 0)
    want = []
    for t in range(P.shape[0]):
        l, qz, qy = orc.run_simple_max_likelihood(
            g["presence"][t], g["ll_cs"], g["ll_ss"], g["cell_idx_ss"], g["snv_idx_cs"], n, m, K,
            g["q_y_init"], g["cell_ptr"], g["snv_ptr"])
        want.append(l)
        gz, gy = e.get_fit(t)
        np.testing.assert_allclose(gz, qz, rtol=Q_RTOL, atol=Q_ATOL)
        assert np.array_equal(gy, qy)
    want = np.array(want)
    np.testing.assert_allclose(lik, want, rtol=ELBO_RTOL)
    assert np.array_equal(np.argsort(-lik, kind="stable"), np.argsort(-want, kind="stable"))
    e.close()


def test_round_pipeline_knobs_do_not_change_bits(eng_mod):
    """The host enqueues rounds ahead of the device (lookahead), sizes the grids by an upper bound on the active
    trees and cuts small sweep launches into shares (split); none of it may change a single bit of any result
    (ELBOs, iteration counts, posteriors), whatever the number of trees in flight."""
    from arborist_b200.arborist import add_normal_clone, prepare_trees
    g = load_golden("small")
    e, _, _ = _coo_engine(eng_mod, g)
    e.set_prior(g["q_y_init"])
    trees, _, _ = golden_frames(g)
    c2i, _, parents, _, normal = prepare_trees(add_normal_clone(trees))
    K = len(c2i)
    kw = dict(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0, keep_posteriors="all")
    e.set_option("split", 1)
    ref = e.fit(**kw)
    ref_q = [e.get_fit(t) for t in range(len(parents))]
    for split, lookahead, slots in ((0, 0, 0), (2, 1, 0), (3, 5, 0), (7, 2, 5), (16, 3, 3), (0, 1, 1)):
        e.set_option("split", split)
        e.set_option("lookahead", lookahead)
        e.set_option("slots", slots)
        got = e.fit(**kw)
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]), (split, lookahead, slots)
        for t in (0, len(parents) - 1):
            qz, qy = e.get_fit(t)
            assert np.array_equal(qz, ref_q[t][0]) and np.array_equal(qy, ref_q[t][1])
        pf = e.last_fit_profile()
        assert pf["idle_rounds"] <= (lookahead or 3) + 1         # rounds enqueued after the last tree finished
    e.close()


def test_device_factorisation_matches_pandas(eng_mod):
    """arb_ingest_ids (ids factorised on the device) vs pd.factorize + arb_ingest_coo: same id order, same
    matrix, and a table with string ids (host path) gives the same ELBOs as its integer twin."""
    import pandas as pd
    from arborist_b200 import arborist
    rng = np.random.default_rng(7)
    n = 20000
    cell = rng.integers(-50, 400, n) * 1000003          # negative and large ids, many repeats
    snv = rng.integers(0, 900, n) * 7 + 10**12
    key = np.unique(np.stack([cell, snv], 1), axis=0)
    rng.shuffle(key)
    cell, snv = key[:, 0].copy(), key[:, 1].copy()
    total = 1 + rng.poisson(0.3, len(cell))
    alt = rng.binomial(total, 0.3)
    e = eng_mod.Engine()
    cu, su = e.ingest_ids(cell, snv, alt, total, 0.001)
    cc, cu_ref = pd.factorize(cell)
    sc, su_ref = pd.factorize(snv)
    assert np.array_equal(cu, cu_ref) and np.array_equal(su, su_ref)
    ptr_a, idx_a, ll_a = e.export_sparse()
    orders_a = e.orders()
    e2 = eng_mod.Engine()
    e2.ingest_coo(cc, sc, alt, total, len(cu_ref), len(su_ref), 0.001)
    ptr_b, idx_b, ll_b = e2.export_sparse()
    assert np.array_equal(ptr_a, ptr_b) and np.array_equal(idx_a, idx_b) and np.array_equal(ll_a, ll_b)
    assert all(np.array_equal(a, b) for a, b in zip(orders_a, e2.orders()))
    e.close(); e2.close()
    # API level: integer ids (device path) vs the same table with string ids (host path)
    g = load_golden("small")
    trees, rc, cl = golden_frames(g)
    lik_i, best_i, _ = arborist(trees, rc, cl, max_iter=25, add_normal=True)
    rc_s = rc.assign(cell=rc["cell"].map(lambda v: f"c{v}"), snv=rc["snv"].map(lambda v: f"s{v}"))
    cl_s = cl.copy()
    cl_s[cl_s.columns[0]] = cl_s[cl_s.columns[0]].map(lambda v: f"s{v}")
    lik_s, best_s, _ = arborist(trees, rc_s, cl_s, max_iter=25, add_normal=True)
    assert [lik_i[t] for t in range(len(trees))] == [lik_s[t] for t in range(len(trees))]
    assert best_i.tree_idx == best_s.tree_idx
    assert np.array_equal(best_i.q_z, best_s.q_z)
    assert [f"c{c}" for c in best_i.cell_to_idx] == list(best_s.cell_to_idx)


def test_stream_word_width_does_not_change_bits(eng_mod):
    """16-bit stream words {row | code << 11} (used when they fit) and the 32-bit form drive the same arithmetic."""
    from arborist_b200.arborist import add_normal_clone, prepare_trees
    g = load_golden("small")
    e, _, _ = _coo_engine(eng_mod, g)
    e.set_prior(g["q_y_init"])
    trees, _, _ = golden_frames(g)
    c2i, _, parents, _, normal = prepare_trees(add_normal_clone(trees))
    K = len(c2i)
    ref = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0, keep_posteriors="all")
    assert e.last_fit_profile()["w16"] == 1
    ref_q = e.get_fit(2)
    e.set_option("w32", 1)
    got = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0, keep_posteriors="all")
    assert e.last_fit_profile()["w16"] == 0
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])
    qz, qy = e.get_fit(2)
    assert np.array_equal(qz, ref_q[0]) and np.array_equal(qy, ref_q[1])
    # the exponent-peeling variants of the sweep (used when a layout cannot prove that its running products
    # stay inside the double range) with both word widths: same iteration counts, ELBOs within rounding
    for w32 in (0, 1):
        e.set_option("w32", w32)
        e.set_option("no_fold", 1)
        got = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0, keep_posteriors="all")
        pf = e.last_fit_profile()
        assert pf["fold_z"] == 0 and pf["fold_y"] == 0 and pf["w16"] == 1 - w32
        np.testing.assert_allclose(got[0], ref[0], rtol=1e-12)
        assert np.array_equal(got[1], ref[1])
        qz, qy = e.get_fit(2)
        np.testing.assert_allclose(qz, ref_q[0], rtol=1e-9, atol=1e-300)
        np.testing.assert_allclose(qy, ref_q[1], rtol=1e-9, atol=1e-300)
    e.close()


def test_slot_count_and_duplicates_do_not_change_bits(eng_mod):
    """Determinism: results are independent of how many trees are in flight, and duplicate
    trees get bit-identical ELBOs (the reference has both properties)."""
    from arborist_b200.arborist import add_normal_clone, prepare_trees
    g = load_golden("small")
    e, _, _ = _coo_engine(eng_mod, g)
    e.set_prior(g["q_y_init"])
    trees, _, _ = golden_frames(g)
    trees = add_normal_clone(trees)
    trees = trees + [trees[3], trees[0]]
    c2i, _, parents, _, normal = prepare_trees(trees)
    K = len(c2i)
    ref = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0)
    assert ref[0][3] == ref[0][-2] and ref[0][0] == ref[0][-1]
    for slots in (1, 3, 5):
        e.set_option("slots", slots)
        got = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0)
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])
    e.set_option("slots", 0)
    for kb in (8, 32):                      # many small panels: same ranking, ELBO within rounding
        e.set_option("panel_kb", kb)
        got = e.fit(parents=parents, n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0)
        np.testing.assert_allclose(got[0], ref[0], rtol=1e-12)
        assert np.array_equal(got[1], ref[1])
    e.close()


def test_against_oracle_on_fresh_inputs(eng_mod):
    """Seeded inputs that are in no fixture: GPU vs the CPU oracle, incl. max_iter=0/1 and a
    tight tolerance (runs to the iteration cap)."""
    from arborist_b200 import arborist, synth
    from oracle import oracle as orc
    cfg = synth.SynthConfig("fresh", 220, 640, 9, 7, seed=4242, coverage=0.06, offtree_frac=0.1,
                            extra_depth=0.6)
    d = synth.generate(cfg)
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    for max_iter, tol in ((0, 1.0), (1, 1.0), (12, 1e-9), (25, 1.0)):
        lik, best, fits = arborist(d.trees, rc, cl, max_iter=max_iter, tolerance=tol, add_normal=True)
        olik, obest, ofits, _ = orc.arborist_oracle(d.trees, rc, cl, max_iter=max_iter, tolerance=tol,
                                                    add_normal=True)
        assert best.tree_idx == obest
        for t in range(len(d.trees)):
            assert lik[t] == pytest.approx(olik[t], rel=ELBO_RTOL)
            assert arborist.last_stats["iters"][t] == ofits[t][3]
            np.testing.assert_allclose(fits[t].q_z, ofits[t][1], rtol=Q_RTOL, atol=Q_ATOL)
            np.testing.assert_allclose(fits[t].q_y, ofits[t][2], rtol=Q_RTOL, atol=Q_ATOL)
            assert_map_equal(fits[t].q_z, ofits[t][1], where="fresh")
            assert_map_equal(fits[t].q_y, ofits[t][2], where="fresh")


def test_device_csv_reader_matches_dataframe_path(eng_mod, tmp_path):
    """arb_csv_* (the file's bytes parsed, factorised and ingested on the device) vs arb_ingest_ids on the DataFrame
    pandas reads from the same file: same id order and the same matrix, for an extra index column and shuffled
    column order, CRLF line ends, a missing final newline and chunks far smaller than the file; tables the device
    parser declines (text ids) go through the host reader with the same result."""
    import pandas as pd
    from arborist_b200 import arborist
    rng = np.random.default_rng(11)
    n = 30000
    key = np.unique(np.stack([rng.integers(-40, 500, n) * 1000003, rng.integers(0, 700, n) * 7 + 10**12], 1), axis=0)
    rng.shuffle(key)
    total = 1 + rng.poisson(0.4, len(key))
    rc = pd.DataFrame({"total": total, "cell": key[:, 0], "note": "x", "alt": rng.binomial(total, 0.3), "snv": key[:, 1]})
    ref = eng_mod.Engine()
    cu_ref, su_ref = ref.ingest_ids(rc["cell"], rc["snv"], rc["alt"], rc["total"], 0.001)
    sparse_ref, orders_ref = ref.export_sparse(), ref.orders()
    ref.close()
    variants = {"plain.csv": dict(index=True), "crlf.csv": dict(index=False, lineterminator="\r\n"),
                "nonl.csv": dict(index=False)}
    for name, kw in variants.items():
        path = tmp_path / name
        rc.to_csv(path, **kw)
        if name == "nonl.csv":
            path.write_bytes(path.read_bytes().rstrip(b"\n"))
        for chunk in (256 << 20, 4096):
            e = eng_mod.Engine()
            cu, su, rows = e.ingest_csv(str(path), 0.001, chunk_bytes=chunk)
            assert rows == len(rc) and np.array_equal(cu, cu_ref) and np.array_equal(su, su_ref), (name, chunk)
            assert all(np.array_equal(a, b) for a, b in zip(e.export_sparse(), sparse_ref)), (name, chunk)
            assert all(np.array_equal(a, b) for a, b in zip(e.orders(), orders_ref))
            e.close()
    # what the device parser declines
    bad = tmp_path / "text.csv"
    rc.assign(cell=rc["cell"].map(lambda v: f"c{v}")).to_csv(bad, index=False)
    e = eng_mod.Engine()
    with pytest.raises(eng_mod.ArboristError) as err:
        e.ingest_csv(str(bad), 0.001)
    assert err.value.status == -5          # ARB_ERR_DECLINED: its own code, so real failures do not re-parse
    e.close()
    with pytest.raises(ValueError):
        (tmp_path / "nohdr.csv").write_text("a,b,c\n1,2,3\n")
        eng_mod.Engine().ingest_csv(str(tmp_path / "nohdr.csv"), 0.001)
    # API level: a path instead of a DataFrame (integer file: device reader; text ids: host reader)
    g = load_golden("small")
    trees, frame, cl = golden_frames(g)
    frame.to_csv(tmp_path / "small.csv", index=False)
    lik_d, best_d, _ = arborist(trees, frame, cl, max_iter=25, add_normal=True)
    lik_f, best_f, _ = arborist(trees, str(tmp_path / "small.csv"), cl, max_iter=25, add_normal=True)
    assert [lik_d[t] for t in range(len(trees))] == [lik_f[t] for t in range(len(trees))]
    assert np.array_equal(best_d.q_z, best_f.q_z) and list(best_d.cell_to_idx) == list(best_f.cell_to_idx)
    frame.assign(cell=frame["cell"].map(lambda v: f"c{v}")).to_csv(tmp_path / "small_text.csv", index=False)
    lik_t, best_t, _ = arborist(trees, str(tmp_path / "small_text.csv"), cl, max_iter=25, add_normal=True)
    assert [lik_d[t] for t in range(len(trees))] == [lik_t[t] for t in range(len(trees))]
    assert [f"c{c}" for c in best_d.cell_to_idx] == list(best_t.cell_to_idx)


@pytest.mark.parametrize("n_clones", [3, 4, 10, 13, 14, 17, 20, 21, 25, 29, 32])
def test_clone_counts_against_oracle(eng_mod, n_clones):
    """Every lanes-per-row width of the seeding / finishing kernels (ceil(K/4) = 1..8 lanes, a power of two or
    not) and the column packing that goes with it, against the CPU oracle on a small seeded workload."""
    from arborist_b200 import arborist, synth
    from oracle import oracle as orc
    cfg = synth.SynthConfig(f"k{n_clones}", 150, 400, n_clones, 5, seed=7000 + n_clones, coverage=0.08,
                            min_proportion=0.01)
    d = synth.generate(cfg)
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    lik, best, fits = arborist(d.trees, rc, cl, max_iter=6, tolerance=1.0, add_normal=True)
    olik, obest, ofits, _ = orc.arborist_oracle(d.trees, rc, cl, max_iter=6, tolerance=1.0, add_normal=True)
    # With three clones the five candidates repeat a topology, so their ELBOs tie to the last bit and the
    # oracle's own argmax depends on the order its OpenMP reduction happened to combine in: the winner has to
    # be one of the oracle's maximisers, and THE oracle's winner whenever that one is unique.
    ties = [t for t in olik if abs(olik[t] - olik[obest]) <= 1e-12 * abs(olik[obest])]
    assert best.tree_idx in ties
    if len(ties) == 1:
        assert best.tree_idx == obest
    for t in range(len(d.trees)):
        assert lik[t] == pytest.approx(olik[t], rel=ELBO_RTOL)
        assert arborist.last_stats["iters"][t] == ofits[t][3]
        np.testing.assert_allclose(fits[t].q_z, ofits[t][1], rtol=Q_RTOL, atol=Q_ATOL)
        np.testing.assert_allclose(fits[t].q_y, ofits[t][2], rtol=Q_RTOL, atol=Q_ATOL)
        assert_map_equal(fits[t].q_z, ofits[t][1], where=f"k{n_clones}")
        assert_map_equal(fits[t].q_y, ofits[t][2], where=f"k{n_clones}")


def test_non_forest_presence_and_errors(eng_mod):
    """Inputs that are not forests go through explicit genotype matrices; bad input is an error,
    not a fallback."""
    from arborist_b200 import arborist, synth
    from oracle import oracle as orc
    d = synth.generate(synth.CONFIGS["tiny"])
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    dag = [[(0, 1), (1, 2), (1, 3), (2, 4), (3, 4)], [(0, 1), (1, 2), (2, 3), (3, 4)]]   # node 4 has two parents
    lik, best, fits = arborist(dag, rc, cl, max_iter=5)
    olik, obest, ofits, _ = orc.arborist_oracle(dag, rc, cl, max_iter=5)
    assert best.tree_idx == obest
    for t in range(2):
        assert lik[t] == pytest.approx(olik[t], rel=ELBO_RTOL)
    with pytest.raises(ValueError, match="same set of clones"):
        arborist([[(0, 1), (1, 2)], [(0, 1), (1, 3)]], rc, cl)
    e = eng_mod.Engine(0)
    with pytest.raises(eng_mod.ArboristError):
        e.dims()
    with pytest.raises(eng_mod.ArboristError):
        eng_mod.Engine(9999)
    e.close()


def test_config3_properties_at_full_shape(eng_mod):
    """Size-independent properties at the headline shape (10k x 20k, 12 clones), 24 trees:
    rows of q sum to 1, the reported ELBO is the maximum of the trace, duplicate trees agree
    bit for bit, and re-running the same shard reproduces every bit."""
    from arborist_b200 import arborist, synth
    d = synth.generate(synth.CONFIGS["config3"])
    trees = d.trees[:22] + [d.trees[5], d.trees[0]]
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    lik, best, fits = arborist(trees, rc, cl, max_iter=25, add_normal=True)
    eng = arborist.last_stats["engine"]()
    assert lik[5] == lik[22] and lik[0] == lik[23]
    assert np.array_equal(fits[5].q_z, fits[22].q_z)
    np.testing.assert_allclose(best.q_z.sum(1), 1.0, rtol=0, atol=1e-12)
    np.testing.assert_allclose(best.q_y.sum(1), 1.0, rtol=0, atol=1e-12)
    lik2, best2, _ = arborist(trees, rc, cl, max_iter=25, add_normal=True)
    assert lik2 == lik and best2.tree_idx == best.tree_idx
    g = load_golden("config3_head")
    for k, t in enumerate(int(i) for i in g["tree_subset"]):
        if t < 22:
            assert lik[t] == pytest.approx(float(g["elbo"][k]), rel=ELBO_RTOL)
    del eng


def test_cli_outputs_match_reference(eng_mod, tmp_path, capsys):
    """CLI level: the command line of the reference's main() (A.py:1133-1191)
    replayed through arborist_b200.arborist.main on the same input files; every output file is diffed
    against what the unmodified reference wrote (tests/golden/cli_tiny/expected, oracle/gen_cli_golden.py)."""
    import pandas as pd
    sys_path = os.path.join(os.path.dirname(__file__), "..", "oracle")
    import sys
    sys.path.insert(0, sys_path)
    try:
        import gen_cli_golden as gc
    finally:
        sys.path.remove(sys_path)
    from arborist_b200.arborist import main
    indir = os.path.join(os.path.dirname(__file__), "golden", "cli_tiny")
    exp = os.path.join(indir, "expected")
    main(gc.command_line(indir, str(tmp_path)))
    out = capsys.readouterr().out
    assert "Arborist complete!" in out and "Tree 4" in out and "ELBO: -345.43" in out
    got, want = pd.read_csv(tmp_path / "ranking.csv"), pd.read_csv(os.path.join(exp, "ranking.csv"))
    assert list(got.columns) == list(want.columns) and list(got["tree_idx"]) == list(want["tree_idx"])
    np.testing.assert_allclose(got["elbo"], want["elbo"], rtol=ELBO_RTOL)
    for name in ("cell_assign.csv", "snv_assign.csv", "tree.txt"):          # discrete outputs: byte-identical
        assert open(tmp_path / name).read() == open(os.path.join(exp, name)).read(), name
    for name in ("q_z.csv", "q_y.csv"):
        g, w = pd.read_csv(tmp_path / name), pd.read_csv(os.path.join(exp, name))
        assert list(g.columns) == list(w.columns) and list(g["id"]) == list(w["id"])
        np.testing.assert_allclose(g.iloc[:, 1:].to_numpy(), w.iloc[:, 1:].to_numpy(), rtol=Q_RTOL, atol=Q_ATOL)


def test_two_clone_candidate_set(eng_mod):
    """K = 2 (the normal clone plus one cluster): (1 - gamma) / (k - 1) is guarded by ``k > 1`` in the reference
    (A.py:552), so the prior row is [gamma] -- mass 0.7, not 1.  The reference's log-domain kernels take it as it
    is: ELBO_0 is inflated by that mass and stays the best, so the returned q(y) is the prior itself.  The first
    pass of such priors runs k_first_pass_general; through arborist() against the oracle."""
    from arborist_b200 import arborist, synth
    from oracle import oracle as orc
    d = synth.generate(synth.SynthConfig("k2", 120, 300, 2, 1, seed=7002, coverage=0.1, min_proportion=0.2))
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    trees = [[(0, 1)]]
    lik, best, fits = arborist(trees, rc, cl, max_iter=5, tolerance=1.0)
    olik, obest, ofits, _ = orc.arborist_oracle(trees, rc, cl, max_iter=5, tolerance=1.0)
    assert best.tree_idx == obest == 0
    assert lik[0] == pytest.approx(olik[0], rel=ELBO_RTOL)
    assert arborist.last_stats["iters"][0] == ofits[0][3]
    np.testing.assert_allclose(fits[0].q_z, ofits[0][1], rtol=Q_RTOL, atol=Q_ATOL)
    np.testing.assert_allclose(fits[0].q_y, ofits[0][2], rtol=Q_RTOL, atol=Q_ATOL)
    assert_map_equal(fits[0].q_z, ofits[0][1], where="k2")


def test_unnormalised_prior_through_the_seam(eng_mod):
    """A q_y_init whose rows do not sum to one, handed to the 13-argument seam (A.py:679-816 takes any matrix):
    first pass with the rows' own mass (k_first_pass_general), then the ordinary sweeps; against the oracle's
    log-domain restatement, for a prior that wins at iteration 0 and one that does not, next to a normalised
    prior on the same cached matrix."""
    from arborist_b200 import run_variational_inference
    from oracle import oracle as orc
    g = load_golden("small")
    n, m, K = int(g["n_cells"]), int(g["n_snvs"]), int(g["n_clones"])
    rng = np.random.default_rng(3)
    args = (g["ll_cs"], g["ll_ss"], g["cell_idx_ss"], g["snv_idx_cs"], n, m, K)
    tail = (g["cell_ptr"], g["snv_ptr"], 25, 1.0)
    for scale in (rng.uniform(0.5, 1.5, size=(m, 1)), np.full((m, 1), 1.3), np.ones((m, 1))):
        q0 = np.ascontiguousarray(g["q_y_init"] * scale)
        for t in (0, 4):
            want = orc.run_variational_inference(g["presence"][t], *args, q0, *tail)
            got = run_variational_inference(g["presence"][t], *args, q0, *tail)
            assert got[0] == pytest.approx(want[0], rel=ELBO_RTOL)
            np.testing.assert_allclose(got[1], want[1], rtol=Q_RTOL, atol=Q_ATOL)
            np.testing.assert_allclose(got[2], want[2], rtol=Q_RTOL, atol=Q_ATOL)


def test_multi_context_fit_is_bit_identical(eng_mod):
    """arb_fit_trees_multi: the candidate set spread over several contexts by one process (tree t -> context
    t mod G, one host thread each) gives the bits of a single-context fit -- ELBOs, iteration counts, traces,
    posteriors and MAP.  The box has one GPU, so the contexts share device 0; on a multi-GPU box arborist()
    picks one context per GPU (``devices=``)."""
    from arborist_b200 import arborist
    from arborist_b200.arborist import add_normal_clone, ingest_read_counts, prepare_trees
    g = load_golden("small")
    trees, rc, cl = golden_frames(g)
    c2i, k2i, parents, _, normal = prepare_trees(add_normal_clone(trees))
    K, T = len(c2i), len(parents)
    kw = dict(n_clones=K, normal_idx=normal, max_iter=25, tolerance=1.0, keep_posteriors="all", trace=True)
    engines = [eng_mod.Engine(0) for _ in range(3)]
    for e in engines:
        ingest_read_counts(e, rc, cl, k2i, float(g["alpha"]), float(g["gamma"]))
    ref = engines[0].fit(parents=parents, **kw)
    ref_q = [engines[0].get_fit(t) for t in range(T)]
    ref_map = [engines[0].get_map(t) for t in range(T)]
    for G in (2, 3):
        got = eng_mod.fit_multi(engines[:G], parents, K, normal_idx=normal, max_iter=25, tolerance=1.0,
                                keep_posteriors="all", trace=True)
        assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1])
        assert np.array_equal(np.nan_to_num(got[2], nan=-1.0), np.nan_to_num(ref[2], nan=-1.0))
        for t in range(T):
            qz, qy = engines[t % G].get_fit(t // G)
            assert np.array_equal(qz, ref_q[t][0]) and np.array_equal(qy, ref_q[t][1])
            z, y = engines[t % G].get_map(t // G)
            assert np.array_equal(z, ref_map[t][0]) and np.array_equal(y, ref_map[t][1])
    for e in engines:
        e.close()
    # API level: devices=[0, 0] (two contexts) against device=0
    lik1, best1, fits1 = arborist(trees, rc, cl, max_iter=25, add_normal=True, device=0)
    lik2, best2, fits2 = arborist(trees, rc, cl, max_iter=25, add_normal=True, devices=[0, 0])
    assert arborist.last_stats["devices"] == [0, 0]
    assert lik1 == lik2 and best1.tree_idx == best2.tree_idx
    for t in range(T):
        assert np.array_equal(fits1[t].q_z, fits2[t].q_z) and np.array_equal(fits1[t].q_y, fits2[t].q_y)
        assert fits1[t].tree == fits2[t].tree == add_normal_clone(trees)[t]
        assert fits1[t].map_assign_y().equals(fits2[t].map_assign_y())
    with pytest.raises(eng_mod.ArboristError):
        e = eng_mod.Engine(0)
        ingest_read_counts(e, rc, cl, k2i, float(g["alpha"]), float(g["gamma"]))
        eng_mod.fit_multi([e, e], parents, K, normal_idx=normal)            # the same context twice


def test_writers_at_scale_use_batched_refits(eng_mod, tmp_path):
    """Row f1: --pickle / q_z of every fit when the posteriors were NOT kept on the device (always the case at
    config 4).  256 config-2-sized fits: the streaming pickle costs ceil(256 / batch) batched re-fits (not 256
    single-tree fits), and its contents equal the eager path (posteriors kept, pd.to_pickle of the dict)."""
    import pandas as pd
    from arborist_b200 import arborist, synth
    cfg = synth.CONFIGS["config2"]
    d = synth.generate(cfg)
    rng = np.random.default_rng(5)
    trees = list(d.trees)
    while len(trees) < 256:
        trees.append(synth._edges(synth.random_tree(rng, cfg.n_clones - 1)))
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    lik_k, best_k, fits_k = arborist(trees, rc, cl, max_iter=25, add_normal=True, device=0, keep_posteriors="all")
    pd.to_pickle(fits_k, tmp_path / "eager.pkl")
    eager = pd.read_pickle(tmp_path / "eager.pkl")
    lik_n, best_n, fits_n = arborist(trees, rc, cl, max_iter=25, add_normal=True, device=0, keep_posteriors="none")
    assert lik_n == lik_k and best_n.tree_idx == best_k.tree_idx
    store = arborist.last_stats["store"]()
    eng = arborist.last_stats["engine"]()
    fits_n.to_pickle(tmp_path / "stream.pkl")
    want_batches = -(-256 // store.batch)
    assert store.fetches == want_batches
    assert eng.last_fit_profile()["refit_batches"] == want_batches <= 2
    back = pd.read_pickle(tmp_path / "stream.pkl")
    assert type(back) is dict and list(back) == list(range(256))
    for t in range(256):
        assert np.array_equal(back[t].q_z, eager[t].q_z) and np.array_equal(back[t].q_y, eager[t].q_y)
        assert back[t].tree == eager[t].tree and back[t].elbo == eager[t].elbo
    # MAP tables of a tree inside the cached batch need no further re-fit
    before = eng.last_fit_profile()["refit_batches"]
    assert fits_n[255].map_assign_z().equals(eager[255].map_assign_z())
    assert eng.last_fit_profile()["refit_batches"] == before


def _with_deep_rows(rc, seed):
    """Give ~4 % of the rows deep coverage: alt = 0 with total 14..80 (w down to -55), every read variant with
    total 85..120 (w up to +880) and a few moderate ones -- the pairs whose log-likelihood ratio leaves
    [ARB_W_LO, ARB_W_HI] and that the reference handles in the log domain without noticing."""
    rng = np.random.default_rng(seed)
    rc = rc.copy()
    idx = rng.choice(len(rc), size=max(12, len(rc) // 25), replace=False)
    third = len(idx) // 3
    tot = rc["total"].to_numpy().copy()
    alt = rc["alt"].to_numpy().copy()
    tot[idx[:third]] = rng.integers(14, 81, third)
    alt[idx[:third]] = 0
    tot[idx[third:2 * third]] = rng.integers(85, 121, third)
    alt[idx[third:2 * third]] = tot[idx[third:2 * third]]
    tot[idx[2 * third:]] = rng.integers(20, 60, len(idx) - 2 * third)
    alt[idx[2 * third:]] = rng.integers(0, 4, len(idx) - 2 * third)
    rc["total"], rc["alt"] = tot, alt
    return rc


def test_deep_coverage_pairs_go_through_the_log_domain(eng_mod):
    """Count pairs whose log-likelihood ratio w = lp - la leaves [-9, 600] (alt = 0 with total >= 14, or >= 83 alt
    reads, at the default alpha) cannot use the product form 1 + expm1(w) S at the reference's accuracy: their
    nonzeros are neutral in the sweeps and the finishing kernels add log(Sbar + S e^w) from side lists, like the
    reference's log-sum-exp (A.py:229-243, 354-369).  Against the oracle: ELBOs, iteration counts, posteriors,
    ranking, MAP -- soft and hard-assignment modes, and the K = 2 first pass (k_first_pass_general)."""
    from arborist_b200 import arborist, synth
    from oracle import oracle as orc
    cfg = synth.SynthConfig("deepcov", 220, 640, 9, 6, seed=4343, coverage=0.06, offtree_frac=0.1, extra_depth=0.6)
    d = synth.generate(cfg)
    rc, cl = _with_deep_rows(d.read_counts_frame(), 1), d.clusters_frame()
    lik, best, fits = arborist(d.trees, rc, cl, max_iter=25, tolerance=1.0, add_normal=True, device=0)
    eng = arborist.last_stats["engine"]()
    lp, la = eng.loglik_table()
    assert ((lp - la) < -9).any() and ((lp - la) > 600).any(), "the workload has no deep pairs"
    olik, obest, ofits, prob = orc.arborist_oracle(d.trees, rc, cl, max_iter=25, tolerance=1.0, add_normal=True)
    assert best.tree_idx == obest
    got = np.array([lik[t] for t in range(len(d.trees))])
    want = np.array([olik[t] for t in range(len(d.trees))])
    np.testing.assert_allclose(got, want, rtol=ELBO_RTOL)
    assert np.array_equal(np.argsort(-got, kind="stable"), np.argsort(-want, kind="stable"))
    for t in range(len(d.trees)):
        assert arborist.last_stats["iters"][t] == ofits[t][3]
        np.testing.assert_allclose(fits[t].q_z, ofits[t][1], rtol=1e-9, atol=Q_ATOL)
        np.testing.assert_allclose(fits[t].q_y, ofits[t][2], rtol=1e-9, atol=Q_ATOL)
        assert_map_equal(fits[t].q_z, ofits[t][1], where="deepcov")
        assert_map_equal(fits[t].q_y, ofits[t][2], where="deepcov")
    # hard-assignment mode on the same matrix (same side lists, HARD variant of the q(z) finishing kernel)
    P = orc.enumerate_presence(orc.add_normal_clone(d.trees)[0], best.clone_to_idx, best.cluster_to_idx)
    K = len(best.clone_to_idx)
    ml, iters = eng.fit(presence=np.ascontiguousarray(P != 0, np.uint8)[None], n_clones=K, max_iter=5, mode="ml",
                        keep_posteriors="all")
    want_ml = orc.run_simple_max_likelihood(P, prob.ll_cs, prob.ll_ss, prob.cell_idx_ss, prob.snv_idx_cs, prob.n_cells,
                                            prob.n_snvs, K, prob.q_y_init, prob.cell_ptr, prob.snv_ptr)
    assert ml[0] == pytest.approx(want_ml[0], rel=ELBO_RTOL)
    # K = 2: prior rows of mass gamma AND deep pairs in the general first pass
    d2 = synth.generate(synth.SynthConfig("k2deep", 120, 300, 2, 1, seed=7003, coverage=0.1, min_proportion=0.2))
    rc2, cl2 = _with_deep_rows(d2.read_counts_frame(), 2), d2.clusters_frame()
    lik2, best2, fits2 = arborist([[(0, 1)]], rc2, cl2, max_iter=5, device=0)
    olik2, _, ofits2, _ = orc.arborist_oracle([[(0, 1)]], rc2, cl2, max_iter=5)
    assert lik2[0] == pytest.approx(olik2[0], rel=ELBO_RTOL)
    np.testing.assert_allclose(fits2[0].q_z, ofits2[0][1], rtol=1e-9, atol=Q_ATOL)


def test_impossible_counts_are_refused(eng_mod):
    """alt > total has no binomial likelihood (scipy gives NaN / -inf): an error that names the pair, never numbers."""
    from arborist_b200 import arborist, synth
    d = synth.generate(synth.CONFIGS["tiny"])
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    bad = rc.copy()
    bad.loc[bad.index[3], ["alt", "total"]] = [5, 2]
    with pytest.raises(eng_mod.ArboristError) as err:
        arborist(d.trees, bad, cl, max_iter=3, add_normal=True, device=0)
    assert err.value.status == -1 and "alt 5, total 2" in str(err.value)


def test_zz_map_tie_rows_are_a_handful(eng_mod):
    """north_star: MAP assignments bit-exact.  The only rows allowed to differ from the oracle are exact
    mathematical ties (see assert_map_equal); across the whole module they must stay a handful."""
    total = sum(v for k, v in MAP_TIE_ROWS.items() if not k.startswith(("ml_", "golden_")))
    print("MAP rows decided by a tie, per test:", {k: v for k, v in MAP_TIE_ROWS.items() if v}, "total", total)
    assert total <= 8, MAP_TIE_ROWS
