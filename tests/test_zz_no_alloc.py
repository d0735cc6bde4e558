"""``arb_base_score_static``: the legacy tree scorer for hosts that must not allocate device memory after set-up
(their device memory belongs to another allocator or is capped by a quota, or their scoring loop cannot absorb the
device-wide stall of a ``cudaMalloc``).  Its temporaries live in one fixed array inside the library image instead
of buffers from the pool, so the checks here are the ones that layout can get wrong: the same scores as
``arb_base_score`` (within this suite's tolerance for the legacy scorer), on two calls in a row that share the array
with different shapes, and a clear refusal beyond the array's size.

Collected last on purpose (``pytest -x`` stops at the first failure); it needs nothing from the tests before it and
passes alone in a fresh process as well."""
import ctypes

import numpy as np
import pytest

RTOL = 1e-9            # tests/test_base_scorer.py uses the same bound for scores of the legacy scorer
STATIC_BYTES = 1 << 20  # ARB_BASE_STATIC_BYTES in include/arborist_b200.h


def _table_and_masks(name):
    """A legacy-scorer table on the device plus the bit masks of its candidate trees, built the way rank_trees does."""
    from arborist_b200 import arborist_base as ab, synth
    trees, rc = synth.legacy_inputs(name)
    tab = ab._Table(rc, 0.001)
    pairs = [tab.path_masks(*ab._clone_paths(tree)) for tree in trees]
    anc = np.ascontiguousarray(np.stack([p[0] for p in pairs]), dtype=np.uint32)
    tmask = np.array([p[1] for p in pairs], np.uint32)
    return tab, anc, tmask


def _score(lib, name, handle, anc, tmask):
    """One scorer call straight through the C ABI (no Engine method in between)."""
    lik = np.full(anc.shape[0], np.nan)
    rc = getattr(lib, name)(handle, anc.shape[0], anc.shape[1], anc, tmask, lik)
    assert rc == 0, lib.arb_last_error().decode()
    return lik


@pytest.mark.gpu
def test_static_scratch_scorer_matches_the_allocating_one_on_consecutive_calls():
    from arborist_b200 import engine
    lib = engine.load_library()
    tab, anc, tmask = _table_and_masks("config2")            # 1,000 cells, 100 trees of 8 clones
    try:
        h = tab.engine._h
        want = _score(lib, "arb_base_score", h, anc, tmask)
        assert np.isfinite(want).all() and len(set(want.tolist())) > 1
        # call 1: every tree, one CTA or a few per tree
        got = _score(lib, "arb_base_score_static", h, anc, tmask)
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=0)
        # call 2, straight after, on the array call 1 left behind: three trees in another order, so every section of
        # the array sits at a different offset and each tree is split over many CTAs (tickets + partial sums in use)
        pick = [7, 0, 42]
        got2 = _score(lib, "arb_base_score_static", h, np.ascontiguousarray(anc[pick]), tmask[pick])
        np.testing.assert_allclose(got2, want[pick], rtol=RTOL, atol=0)
        # call 3: back to the full set; what call 2 left behind must not leak into it, and the bits repeat
        got3 = _score(lib, "arb_base_score_static", h, anc, tmask)
        assert (got3 == got).all()
        assert got.argmax() == want.argmax() and (np.argsort(-got, kind="stable") == np.argsort(-want, kind="stable")).all()
        # the allocating entry point is untouched by all this
        assert (_score(lib, "arb_base_score", h, anc, tmask) == want).all()
    finally:
        tab.close()


@pytest.mark.gpu
def test_static_scratch_scorer_with_the_widest_split():
    """20,000 cells and two trees: 64 CTAs per tree leave their partial sums in the array and the last one to arrive
    adds them up.  Repeated calls must give the same bits whichever CTA happens to be last."""
    from arborist_b200 import engine
    lib = engine.load_library()
    rng = np.random.default_rng(7)
    n_rows, n_cells, n_clusters, K = 400_000, 20_000, 6, 5
    cell = rng.integers(0, n_cells, n_rows).astype(np.int32)
    cluster = rng.integers(0, n_clusters, n_rows).astype(np.int32)
    total = rng.integers(1, 6, n_rows).astype(np.int32)
    alt = rng.binomial(total, 0.3).astype(np.int32)
    eng = engine.Engine(0)
    try:
        eng.base_ingest(cell, cluster, total, alt, n_cells, n_clusters, 0.001)
        anc = rng.integers(1, 1 << n_clusters, (2, K)).astype(np.uint32)
        tmask = np.array([(1 << n_clusters) - 1, 0b011011], np.uint32)
        want = _score(lib, "arb_base_score", eng._h, anc, tmask)
        first = _score(lib, "arb_base_score_static", eng._h, anc, tmask)
        np.testing.assert_allclose(first, want, rtol=RTOL, atol=0)
        for _ in range(5):
            assert (_score(lib, "arb_base_score_static", eng._h, anc, tmask) == first).all()
        # through the Engine method as well
        assert (eng.base_score_static(anc, tmask) == first).all()
    finally:
        eng.close()


@pytest.mark.gpu
def test_static_scratch_scorer_on_a_second_context_of_the_same_gpu():
    """The array belongs to the GPU, not to a context: a second context finds it set up and gets the same scores."""
    from arborist_b200 import engine
    lib = engine.load_library()
    tab_a, anc, tmask = _table_and_masks("tiny")
    tab_b, _, _ = _table_and_masks("tiny")
    try:
        want = _score(lib, "arb_base_score", tab_a.engine._h, anc, tmask)
        got_a = _score(lib, "arb_base_score_static", tab_a.engine._h, anc, tmask)
        got_b = _score(lib, "arb_base_score_static", tab_b.engine._h, anc, tmask)
        np.testing.assert_allclose(got_a, want, rtol=RTOL, atol=0)
        assert (got_a == got_b).all()
    finally:
        tab_a.close()
        tab_b.close()


@pytest.mark.gpu
def test_static_scratch_scorer_refuses_what_does_not_fit():
    from arborist_b200 import engine
    lib = engine.load_library()
    tab, anc, tmask = _table_and_masks("tiny")
    try:
        h = tab.engine._h
        K = anc.shape[1]
        n_big = STATIC_BYTES // (4 * K) + 1                   # the staged ancestor masks alone exceed the array
        big_anc = np.ascontiguousarray(np.resize(anc, (n_big, K)))
        big_tm = np.resize(tmask, n_big)
        lik = np.full(n_big, np.nan)
        assert lib.arb_base_score_static(h, n_big, K, big_anc, big_tm, lik) == -3      # ARB_ERR_UNSUPPORTED
        msg = lib.arb_last_error().decode()
        assert str(STATIC_BYTES) in msg and "smaller groups" in msg
        assert np.isnan(lik).all()                            # refused before anything ran
        # the refusal leaves the array usable, and the allocating entry point still takes the big set
        np.testing.assert_allclose(_score(lib, "arb_base_score_static", h, anc, tmask),
                                   _score(lib, "arb_base_score", h, anc, tmask), rtol=RTOL, atol=0)
        assert np.isfinite(_score(lib, "arb_base_score", h, big_anc, big_tm)).all()
        # argument errors are statuses, not crashes
        assert lib.arb_base_score_static(h, 0, K, anc, tmask, lik) == -1
        raw = ctypes.CDLL(engine.library_path())
        raw.arb_base_score_static.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p,
                                              ctypes.c_void_p, ctypes.c_void_p]
        assert raw.arb_base_score_static(None, 1, K, None, None, None) == -1
    finally:
        tab.close()
