"""Generate tests/golden/*.npz from the UNMODIFIED reference (runs only in the
build container, where /root/reference exists; the fixtures travel, this
script's import of the reference does not).

    python oracle/gen_golden.py [names...]      # default: all

What is recorded, per workload (inputs come from arborist_b200.synth with the
seed stored in the file, or are stored verbatim for the example stand-in):

* API level   -- ``arborist(...)`` (A.py:906): ELBO per tree, argmax tree, MAP
                 cell/SNV labels, and (small cases) the full q_z / q_y.
* driver level -- iteration count and ELBO trace per tree, from a Python
                 re-statement of A.py:733-816 that calls the reference's own
                 @njit kernels; its final ELBO is asserted bit-equal to
                 ``run_variational_inference`` before anything is written.
* kernel level -- compute_q_z_sparse / compute_q_y_sparse /
                 compute_likelihood_sparse / entropy / KL on explicit arrays.
* ingest level -- cell/SNV index order, CSR/CSC arrays, prior, presence.

numba thread count is recorded: the reference's ELBO moves by ~3e-12 relative
with the thread count (prange reduction, A.py:409-419).
"""
from __future__ import annotations

import os
import sys
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")

REF = "/root/reference/src"


def load_reference():
    if not os.path.isdir(REF):
        raise SystemExit("reference not present: goldens can only be generated in the build container")
    sys.modules.setdefault("pygraphviz", types.ModuleType("pygraphviz"))   # drawing only
    sys.path.insert(0, REF)
    import arborist  # noqa: F401

    return sys.modules["arborist.arborist"]


def traced_driver(A, P, prob, max_iter, tol):
    """A.py:733-816 restated around the reference's own jitted kernels."""
    n, m, K = prob["n_cells"], prob["n_snvs"], prob["n_clones"]
    cs = (prob["cell_ptr"], prob["snv_idx_cs"], prob["ll_cs"])
    ss = (prob["snv_ptr"], prob["cell_idx_ss"], prob["ll_ss"])
    q0 = prob["q_y_init"]
    lq0 = A.log_array(q0)
    lqy = lq0.copy()
    qz = A.compute_q_z_sparse(*cs, lqy, P, n, K)
    lqz = A.log_array(qz)
    best = (A.compute_likelihood_sparse(*cs, lqy, lqz, P, K, n) + A.compute_entropy(qz)
            - A.compute_kl_divergence(q0, q0))
    trace = [best]
    bz, by = qz.copy(), q0.copy()
    conv = False
    it = 0
    for it in range(max_iter):
        qy = A.compute_q_y_sparse(*ss, lqz, P, m, K, lq0)
        lqy = A.log_array(qy)
        qz = A.compute_q_z_sparse(*cs, lqy, P, n, K)
        lqz = A.log_array(qz)
        lik = A.compute_likelihood_sparse(*cs, lqy, lqz, P, K, n)
        elbo = lik + A.compute_entropy(qz) - A.compute_kl_divergence(qy, q0)
        trace.append(elbo)
        if abs(elbo - best) < tol:
            conv = True
        if elbo > best:
            bz, by, best = qz.copy(), qy.copy(), elbo
        if conv:
            break
    n_it = len(trace) - 1
    return best, bz, by, n_it, np.array(trace)


def prepare_with_reference(A, trees, rc, clusters, alpha, gamma, add_normal):
    """Replays A.py:1004-1069 with the reference's own helpers to expose the arrays."""
    import networkx as nx

    if add_normal:
        trees = A.add_normal_clone(trees)
    tree = trees[0]
    g = nx.DiGraph(tree)
    normal = [x for x in g if g.in_degree(x) == 0][0]
    clones = sorted(A.tree_to_clone_set(tree))
    clusters_l = sorted(c for c in clones if c != normal)
    c2i = {c: i for i, c in enumerate(clones)}
    k2i = {c: i for i, c in enumerate(clusters_l)}
    alt_sum = rc.groupby("snv")["alt"].sum()
    valid = alt_sum[alt_sum > 0].index
    rc = rc[rc["snv"].isin(valid)].copy()
    rc, cell_to_idx, snv_to_idx = A.precompute_log_likelihoods(rc, alpha)
    q0 = A.initialize_q_y(clusters, k2i, snv_to_idx, gamma)
    ci, si, ll = A.build_sparse_input(rc, cell_to_idx, snv_to_idx)
    n, m = len(cell_to_idx), len(snv_to_idx)
    cell_ptr, _, snv_cs, ll_cs = A.build_index_pointers(ci, si, n, log_like_matrix=ll)
    snv_ptr, _, cell_ss, ll_ss = A.build_index_pointers(si, ci, m, log_like_matrix=ll)
    return dict(trees=trees, clone_to_idx=c2i, cluster_to_idx=k2i, cell_to_idx=cell_to_idx,
                snv_to_idx=snv_to_idx, q_y_init=q0, cell_ptr=cell_ptr, snv_idx_cs=snv_cs,
                ll_cs=ll_cs, snv_ptr=snv_ptr, cell_idx_ss=cell_ss, ll_ss=ll_ss,
                n_cells=n, n_snvs=m, n_clones=len(clones), coo_cell=ci, coo_snv=si, coo_ll=ll)


def trees_to_array(trees):
    return np.array([[list(e) for e in t] for t in trees], dtype=np.int64)


def run_case(A, name, trees, rc, clusters, *, alpha=0.001, gamma=0.7, max_iter=25, tol=1.0,
             add_normal=True, keep_q="all", tree_subset=None, extra=None, store_inputs=False,
             store_best=True):
    import numba

    threads = min(os.cpu_count() or 1, 8)
    t0 = time.time()
    if tree_subset is not None:
        trees = [trees[i] for i in tree_subset]
    lik, best_fit, fits = A.arborist(trees, rc.copy(), clusters.copy(), alpha=alpha,
                                     max_iter=max_iter, tolerance=tol, gamma=gamma,
                                     add_normal=add_normal, threads=threads)
    t_api = time.time() - t0
    prob = prepare_with_reference(A, trees, rc.copy(), clusters.copy(), alpha, gamma, add_normal)
    T = len(trees)
    elbo = np.array([lik[i] for i in range(T)])
    iters = np.zeros(T, dtype=np.int32)
    traces = np.full((T, max_iter + 1), np.nan)
    presences = []
    for t in range(T):
        P = A.enumerate_presence(prob["trees"][t], prob["clone_to_idx"], prob["cluster_to_idx"])
        presences.append(P)
        b, bz, by, n_it, tr = traced_driver(A, P, prob, max_iter, tol)
        assert b == elbo[t], (name, t, b, elbo[t])           # bit-equal to the real driver
        assert np.array_equal(bz, fits[t].q_z) and np.array_equal(by, fits[t].q_y)
        iters[t] = n_it
        traces[t, : n_it + 1] = tr
    cells = np.array(list(prob["cell_to_idx"].keys()))
    snvs = np.array(list(prob["snv_to_idx"].keys()))
    map_z = np.stack([fits[t].q_z.argmax(axis=1) for t in range(T)]).astype(np.int16)
    map_y = np.stack([fits[t].q_y.argmax(axis=1) for t in range(T)]).astype(np.int16)
    out = dict(
        name=name, numba_threads=threads, alpha=alpha, gamma=gamma, max_iter=max_iter, tol=tol,
        add_normal=add_normal, elbo=elbo, iters=iters, traces=traces,
        best_idx=best_fit.tree_idx, map_z=map_z, map_y=map_y,
        presence=np.stack(presences), cell_order=cells, snv_order=snvs,
        clones=np.array(list(prob["clone_to_idx"].keys())),
        clusters=np.array(list(prob["cluster_to_idx"].keys())),
        n_cells=prob["n_cells"], n_snvs=prob["n_snvs"], n_clones=prob["n_clones"],
        nnz=prob["coo_cell"].shape[0],
        trees=trees_to_array(trees),            # as given to arborist() (before add_normal)
        ref_versions=f"numba {numba.__version__}",
    )
    if tree_subset is not None:
        out["tree_subset"] = np.array(tree_subset)
    if keep_q == "all":
        out["q_z"] = np.stack([fits[t].q_z for t in range(T)])
        out["q_y"] = np.stack([fits[t].q_y for t in range(T)])
    else:                                       # a slice of rows + the best tree in full
        r = int(keep_q)
        out["q_z_rows"] = np.stack([fits[t].q_z[:r] for t in range(T)])
        out["q_y_rows"] = np.stack([fits[t].q_y[:r] for t in range(T)])
        if store_best:                          # config 4: 24 MB of doubles, left out of the fixture
            out["q_z_best"] = best_fit.q_z
            out["q_y_best"] = best_fit.q_y
    if store_inputs:
        out.update(rc_snv=rc["snv"].to_numpy(), rc_cell=rc["cell"].to_numpy(),
                   rc_alt=rc["alt"].to_numpy(), rc_total=rc["total"].to_numpy(),
                   cl_snv=clusters["snv"].to_numpy(), cl_label=clusters["cluster"].to_numpy())
    if keep_q == "all":
        # ingest + kernel level (small cases only)
        out.update(q_y_init=prob["q_y_init"], cell_ptr=prob["cell_ptr"],
                   snv_idx_cs=prob["snv_idx_cs"], ll_cs=prob["ll_cs"], snv_ptr=prob["snv_ptr"],
                   cell_idx_ss=prob["cell_idx_ss"], ll_ss=prob["ll_ss"],
                   coo_cell=prob["coo_cell"], coo_snv=prob["coo_snv"], coo_ll=prob["coo_ll"])
        P = presences[0]
        n, m, K = prob["n_cells"], prob["n_snvs"], prob["n_clones"]
        rng = np.random.default_rng(7)
        qy = rng.dirichlet(np.ones(K - 1) * 0.3, size=m)
        qy[0, :] = 0.0
        qy[0, 0] = 1.0                                        # exercises log_array's -inf
        lqy = A.log_array(qy)
        kz = A.compute_q_z_sparse(prob["cell_ptr"], prob["snv_idx_cs"], prob["ll_cs"], lqy, P, n, K)
        lkz = A.log_array(kz)
        ky = A.compute_q_y_sparse(prob["snv_ptr"], prob["cell_idx_ss"], prob["ll_ss"], lkz, P, m, K,
                                  A.log_array(prob["q_y_init"]))
        klik = A.compute_likelihood_sparse(prob["cell_ptr"], prob["snv_idx_cs"], prob["ll_cs"],
                                           lqy, lkz, P, K, n)
        out.update(k_qy_in=qy, k_qz=kz, k_qy=ky, k_lik=klik, k_ent=A.compute_entropy(kz),
                   k_kl=A.compute_kl_divergence(ky, prob["q_y_init"]))
        # hard-assignment mode (A.py:598-676), same 13-arg signature
        ml, mlz, mly = A.run_simple_max_likelihood(
            P, prob["ll_cs"], prob["ll_ss"], prob["cell_idx_ss"], prob["snv_idx_cs"], n, m, K,
            prob["q_y_init"], prob["cell_ptr"], prob["snv_ptr"])
        out.update(ml_lik=ml, ml_qz=mlz, ml_qy=mly)
    if extra:
        out.update(extra)
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: T={T} n={prob['n_cells']} m={prob['n_snvs']} K={prob['n_clones']} "
          f"nnz={out['nnz']} api={t_api:.1f}s iters[min/mean/max]={iters.min()}/"
          f"{iters.mean():.1f}/{iters.max()} best={best_fit.tree_idx} -> {path} "
          f"({os.path.getsize(path) / 1e6:.2f} MB)", flush=True)


def example_standin():
    """BASELINE config 1: the bundled trees + input clustering with synthetic counts
    (example/input/read_counts.csv is missing from the checkout)."""
    import pandas as pd

    from arborist_b200.utils import read_trees

    ex = "/root/reference/example"
    trees = read_trees(f"{ex}/input/candidate_trees.txt")
    clusters = pd.read_csv(f"{ex}/input/input_clustering.csv", header=None, names=["snv", "cluster"])
    cz = pd.read_csv(f"{ex}/output/cell_to_clone_labels.csv")
    sy = pd.read_csv(f"{ex}/output/snv_to_cluster_labels.csv")
    rng = np.random.default_rng(1001)
    # ancestor closure under tree 0
    par = {c: p for p, c in trees[0]}
    def anc_or_self(cl, clone):
        x = clone
        while x is not None:
            if x == cl:
                return True
            x = par.get(x)
        return False
    cells, snvs = cz["id"].to_numpy(), sy["id"].to_numpy()
    n, m = len(cells), len(snvs)
    mask = rng.random((m, n)) < 0.02
    js, is_ = np.nonzero(mask)
    total = 1 + rng.poisson(0.1, size=js.shape[0])
    labels_z, labels_y = cz["assignment"].to_numpy(), sy["assignment"].to_numpy()
    closure = {(cl, c): anc_or_self(cl, c) for cl in set(labels_y) for c in set(labels_z)}
    pres = np.fromiter((closure[(labels_y[j], labels_z[i])] for j, i in zip(js, is_)), dtype=bool,
                       count=js.shape[0])
    a = 0.001
    alt = rng.binomial(total, np.where(pres, 0.5 - a + 0.5 * a / 3, a / 3))
    rc = pd.DataFrame({"snv": snvs[js], "cell": cells[is_], "alt": alt, "total": total})
    return trees, rc, clusters


def main(argv):
    from arborist_b200 import synth

    A = load_reference()
    # config5_full (~40 min) and config4_head (~1.5 h, ~20 GB) only when named
    want = set(argv) or {"tiny", "small", "small_nonormal", "k2", "deep30", "config2", "config3_head",
                         "example_standin"}
    if "tiny" in want:
        d = synth.generate(synth.CONFIGS["tiny"])
        run_case(A, "tiny", d.trees, d.read_counts_frame(), d.clusters_frame(), store_inputs=True)
    if "small" in want:
        d = synth.generate(synth.CONFIGS["small"])
        run_case(A, "small", d.trees, d.read_counts_frame(), d.clusters_frame(), store_inputs=True)
    if "small_nonormal" in want:
        # trees that already carry a root clone; API defaults max_iter=10; label 0 quirk
        d = synth.generate(synth.CONFIGS["small"])
        from arborist_b200.arborist import add_normal_clone
        trees = add_normal_clone(d.trees)
        cl = d.clusters_frame()
        run_case(A, "small_nonormal", trees, d.read_counts_frame(), cl, add_normal=False,
                 max_iter=10, gamma=0.9, alpha=0.01, store_inputs=True)
    if "k2" in want:
        # K = 2: one cluster, the prior row is [gamma] (mass 0.7, A.py:552-567) and ELBO_0 stays the best
        cfg = synth.SynthConfig("k2", 120, 300, 2, 1, seed=7002, coverage=0.1, min_proportion=0.2)
        d = synth.generate(cfg)
        run_case(A, "k2", [[(0, 1)]], d.read_counts_frame(), d.clusters_frame(), add_normal=False, max_iter=5,
                 store_inputs=True)
    if "deep30" in want:
        cfg = synth.SynthConfig("deep30", 500, 1000, 30, 10, seed=1105, coverage=0.04,
                                offtree_frac=0.30, deep=True)
        d = synth.generate(cfg)
        run_case(A, "deep30", d.trees, d.read_counts_frame(), d.clusters_frame(), store_inputs=True)
    if "config2" in want:
        d = synth.generate(synth.CONFIGS["config2"])
        run_case(A, "config2", d.trees, d.read_counts_frame(), d.clusters_frame(), keep_q=32)
    if "example_standin" in want:
        trees, rc, clusters = example_standin()
        run_case(A, "example_standin", trees, rc, clusters, add_normal=False, keep_q=32,
                 store_inputs=True)
    if "config3_head" in want:
        d = synth.generate(synth.CONFIGS["config3"])
        run_case(A, "config3_head", d.trees, d.read_counts_frame(), d.clusters_frame(),
                 keep_q=32, tree_subset=[0, 1, 2, 3])
    big_cases(A, want)


def big_cases(A, want):
    """BASELINE configs 5 and 4 at their full shapes (VERDICT r1, row g)."""
    from arborist_b200 import synth

    if "config5_full" in want:
        # 5k x 10k, K = 30 deep trees, 30 % off-tree SNVs: the first 7 trees + the generating tree
        d = synth.generate(synth.CONFIGS["config5"])
        sub = sorted(set(range(7)) | {d.true_tree_index})
        run_case(A, "config5_full", d.trees, d.read_counts_frame(), d.clusters_frame(),
                 keep_q=32, tree_subset=sub, extra=dict(true_tree_index=d.true_tree_index))
    if "config4_head" in want:
        # 100k x 50k, nnz 99 M, K = 20: the first 2 trees, max_iter capped at 2 so the numba
        # reference finishes here ("first 2 trees ... parity sample")
        d = synth.generate(synth.CONFIGS["config4"])
        run_case(A, "config4_head", d.trees, d.read_counts_frame(), d.clusters_frame(),
                 keep_q=32, tree_subset=[0, 1], max_iter=2, store_best=False)


if __name__ == "__main__":
    main(sys.argv[1:])
