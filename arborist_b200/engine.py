"""ctypes binding of ``libarborist_b200.so`` -- the C ABI in ``include/arborist_b200.h``.

``Engine`` is one ``arb_ctx`` (one GPU).  There is no CPU path: a missing library,
a library that fails to load or a machine without a CUDA device raises
``ArboristError`` / ``RuntimeError``; nothing here computes on the host.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _build

__all__ = ["Engine", "ArboristError", "device_count", "library_path", "load_library", "fit_multi"]

_LIB = None

_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_ctx = C.c_void_p
_opt = C.c_void_p           # optional pointer arguments are passed as raw addresses


class ArboristError(RuntimeError):
    """A call into libarborist_b200 returned a negative status."""

    def __init__(self, status, message):
        super().__init__(f"libarborist_b200 error {status}: {message}")
        self.status = status


def library_path() -> str:
    return _build.LIB


_SIGNATURES = {
    "arb_abi_version": (C.c_int, []),
    "arb_last_error": (C.c_char_p, []),
    "arb_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "arb_ctx_create": (C.c_int, [C.c_int, C.POINTER(_ctx)]),
    "arb_ctx_destroy": (C.c_int, [_ctx]),
    "arb_ingest_coo": (C.c_int, [_ctx, C.c_int64, _i32p, _i32p, _i32p, _i32p, C.c_int32, C.c_int32,
                                 C.c_double]),
    "arb_ingest_ids": (C.c_int, [_ctx, C.c_int64, _i64p, _i64p, _i64p, _i64p, C.c_double,
                                 C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "arb_get_id_uniques": (C.c_int, [_ctx, _i64p, _i64p]),
    "arb_ingest_csr": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int64, _i32p, _i32p, _f64p]),
    "arb_get_dims": (C.c_int, [_ctx, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                               C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "arb_get_orders": (C.c_int, [_ctx, _i32p, _i32p]),
    "arb_get_count_pairs": (C.c_int, [_ctx, _i32p, _i32p]),
    "arb_get_loglik_table": (C.c_int, [_ctx, _f64p, _f64p]),
    "arb_set_loglik_table": (C.c_int, [_ctx, C.c_int32, _f64p, _f64p]),
    "arb_export_sparse": (C.c_int, [_ctx, C.c_int, _i32p, _i32p, _f64p]),
    "arb_set_prior_labels": (C.c_int, [_ctx, _i32p, C.c_int32, C.c_double]),
    "arb_set_prior": (C.c_int, [_ctx, _f64p, C.c_int32]),
    "arb_get_prior": (C.c_int, [_ctx, _f64p]),
    "arb_fit_trees": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32, _i32p, C.c_int32, C.c_double,
                                C.c_int32, C.c_int32, _f64p, _i32p, _opt]),
    "arb_fit_presence": (C.c_int, [_ctx, C.c_int32, C.c_int32, _u8p, C.c_int32, C.c_double, C.c_int32,
                                   C.c_int32, _f64p, _i32p, _opt]),
    "arb_posterior_bytes": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int64)]),
    "arb_fit_trees_multi": (C.c_int, [C.POINTER(_ctx), C.c_int32, C.c_int32, C.c_int32, C.c_int32, _i32p, C.c_int32,
                                      C.c_double, C.c_int32, C.c_int32, _f64p, _i32p, _opt]),
    "arb_get_fit": (C.c_int, [_ctx, C.c_int32, _f64p, _f64p]),
    "arb_get_fits": (C.c_int, [_ctx, C.c_int32, C.c_int32, _opt, _opt]),
    "arb_get_map": (C.c_int, [_ctx, C.c_int32, _i32p, _i32p]),
    "arb_get_presence": (C.c_int, [_ctx, C.c_int32, _u8p]),
    "arb_qz_update": (C.c_int, [_ctx, C.c_int32, _u8p, _f64p, _f64p]),
    "arb_qy_update": (C.c_int, [_ctx, C.c_int32, _u8p, _f64p, _f64p]),
    "arb_elbo_terms": (C.c_int, [_ctx, C.c_int32, _u8p, _f64p, _f64p, C.POINTER(C.c_double),
                                 C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "arb_last_fit_stats": (C.c_int, [_ctx, C.POINTER(C.c_int64), C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "arb_last_fit_profile": (C.c_int, [_ctx, _f64p, C.c_int32]),
    "arb_set_option": (C.c_int, [_ctx, C.c_char_p, C.c_int64]),
    "arb_pool_trim": (C.c_int, []),
    "arb_csv_begin": (C.c_int, [_ctx, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64]),
    "arb_csv_feed": (C.c_int, [_ctx, C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    "arb_csv_finish": (C.c_int, [_ctx, C.c_double, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "arb_base_ingest": (C.c_int, [_ctx, C.c_int64, _i32p, _i32p, _i32p, _i32p, C.c_int32, C.c_int32, C.c_double]),
    "arb_base_score": (C.c_int, [_ctx, C.c_int32, C.c_int32, _u32p, _u32p, _f64p]),
    "arb_base_score_static": (C.c_int, [_ctx, C.c_int32, C.c_int32, _u32p, _u32p, _f64p]),
    "arb_base_cells": (C.c_int, [_ctx, C.c_int32, C.c_int32, _u32p, _u32p, _opt, _f64p, _i32p, _f64p, _u8p]),
}


def load_library():
    """Load the CUDA library (building it first only if the sources are newer and nvcc exists)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if _build.is_stale() and _build.nvcc_path() is not None:
        _build.build_library()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing and nvcc is not available to build it; arborist_b200 has no CPU "
            "fallback (run `python -c 'import __graft_entry__ as g; g.build()'` on a machine with nvcc)")
    lib = C.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.arb_abi_version() != 2:
        raise RuntimeError(f"{path}: ABI version {lib.arb_abi_version()} != 2")
    _LIB = lib
    return lib


def _check(status):
    if status != 0:
        raise ArboristError(status, load_library().arb_last_error().decode("utf-8", "replace"))


def trim_pool() -> None:
    """Return the device memory parked in the library's allocation pool to the driver."""
    _check(load_library().arb_pool_trim())


def device_count() -> int:
    n = C.c_int(0)
    _check(load_library().arb_device_count(C.byref(n)))
    return n.value


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Engine:
    """One GPU context: ingest once, set the prior, fit shards of trees, fetch results."""

    def __init__(self, device=None):
        self._lib = load_library()
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
            n = device_count()
            if n > 0:
                device %= n
        self.device = int(device)
        h = _ctx()
        _check(self._lib.arb_ctx_create(self.device, C.byref(h)))
        self._h = h
        self.n_clones = None

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.arb_ctx_destroy(h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- ingest ---------------------------------------------------------------
    def ingest_coo(self, cell_codes, snv_codes, alt, total, n_cell_codes, n_snv_codes, alpha):
        cell_codes, snv_codes = _i32(cell_codes), _i32(snv_codes)
        alt64, tot64 = np.asarray(alt), np.asarray(total)
        for name, v in (("alt", alt64), ("total", tot64)):
            if v.size and (v.max() > np.iinfo(np.int32).max or v.min() < np.iinfo(np.int32).min):
                raise ValueError(f"read-count column '{name}' does not fit int32")
        _check(self._lib.arb_ingest_coo(self._h, cell_codes.shape[0], cell_codes, snv_codes,
                                        _i32(alt64), _i32(tot64), int(n_cell_codes),
                                        int(n_snv_codes), float(alpha)))

    def ingest_ids(self, cell_ids, snv_ids, alt, total, alpha):
        """Long table with int64 ids; the first-appearance factorisation runs on the device.
        Returns ``(cell_uniques, snv_uniques)`` (ids in code order, like ``pd.factorize``)."""
        arrs = [np.ascontiguousarray(a, dtype=np.int64) for a in (cell_ids, snv_ids, alt, total)]
        nc, ns = C.c_int32(), C.c_int32()
        _check(self._lib.arb_ingest_ids(self._h, len(arrs[0]), *arrs, float(alpha), C.byref(nc), C.byref(ns)))
        cu, su = np.empty(nc.value, np.int64), np.empty(ns.value, np.int64)
        _check(self._lib.arb_get_id_uniques(self._h, cu, su))
        return cu, su

    def ingest_csv(self, path, alpha, chunk_bytes=256 << 20):
        """Read-count CSV with integer columns ``cell, snv, alt, total`` (any order, other columns ignored) parsed,
        factorised and ingested on the device.  Returns ``(cell_uniques, snv_uniques, n_rows)``; raises
        ``ArboristError`` with status ``ARB_ERR_DECLINED`` (-5) when the file needs a general CSV reader."""
        size = os.path.getsize(path)
        with open(path, "rb") as f:
            header = f.readline()
            names = [h.strip().strip('"') for h in header.decode("utf-8", "replace").rstrip("\r\n").split(",")]
            try:
                cols = [names.index(c) for c in ("cell", "snv", "alt", "total")]
            except ValueError:
                raise ValueError(f"{path}: the header needs the columns cell, snv, alt, total (found {names})")
            first = f.read(1 << 16)
            per_row = max(4.0, len(first) / max(1, first.count(b"\n")))
            _check(self._lib.arb_csv_begin(self._h, *cols, len(names), int((size - len(header)) / per_row * 1.05) + 16))
            carry, total = first, C.c_int64(0)
            while True:
                block = f.read(chunk_bytes)
                data = carry + block
                if not block:                                   # end of file: the last line may lack its newline
                    if data and not data.endswith(b"\n"):
                        data += b"\n"
                    cut = len(data)
                else:
                    cut = data.rfind(b"\n") + 1
                if cut:
                    _check(self._lib.arb_csv_feed(self._h, data[:cut], cut, C.byref(total)))
                carry = data[cut:]
                if not block:
                    break
        n_rows, nc, ns = C.c_int64(), C.c_int32(), C.c_int32()
        _check(self._lib.arb_csv_finish(self._h, float(alpha), C.byref(n_rows), C.byref(nc), C.byref(ns)))
        cu, su = np.empty(nc.value, np.int64), np.empty(ns.value, np.int64)
        _check(self._lib.arb_get_id_uniques(self._h, cu, su))
        return cu, su, n_rows.value

    # ---- legacy scorer (arb_base.cu; reference arborist_base.py) ---------------------------------
    def base_ingest(self, cell_code, cluster_code, total, alt, n_cells, n_clusters, alpha):
        arrs = [_i32(a) for a in (cell_code, cluster_code, total, alt)]
        _check(self._lib.arb_base_ingest(self._h, len(arrs[0]), *arrs, int(n_cells), int(n_clusters), float(alpha)))
        self._base_cells = int(n_cells)

    def base_score(self, anc_masks, tree_masks):
        anc = np.ascontiguousarray(anc_masks, dtype=np.uint32)
        tm = np.ascontiguousarray(tree_masks, dtype=np.uint32)
        lik = np.empty(anc.shape[0], np.float64)
        _check(self._lib.arb_base_score(self._h, anc.shape[0], anc.shape[1], anc, tm, lik))
        return lik

    def base_score_static(self, anc_masks, tree_masks):
        """``base_score`` without any device allocation (``arb_base_score_static``): for hosts whose device memory
        belongs to another allocator, that run under a memory quota, or that cannot take the driver-wide stall of a
        ``cudaMalloc`` inside their scoring loop.  The temporaries live in a fixed 1 MiB array of the library image,
        so a call that needs more raises ``ArboristError`` (status -3): pass the trees in smaller groups.  One array
        per GPU: do not run two of these calls on the same GPU at the same time."""
        anc = np.ascontiguousarray(anc_masks, dtype=np.uint32)
        tm = np.ascontiguousarray(tree_masks, dtype=np.uint32)
        lik = np.empty(anc.shape[0], np.float64)
        _check(self._lib.arb_base_score_static(self._h, anc.shape[0], anc.shape[1], anc, tm, lik))
        return lik

    def base_cells(self, anc_masks, tree_masks, want_loglik=False):
        """Per-cell detail of the listed trees: ``(loglik or None, posterior, clone_idx, entropy, present)``."""
        anc = np.ascontiguousarray(anc_masks, dtype=np.uint32)
        tm = np.ascontiguousarray(tree_masks, dtype=np.uint32)
        S, K, n = anc.shape[0], anc.shape[1], self._base_cells
        ll = np.empty((S, n, K), np.float64) if want_loglik else None
        post, idx = np.empty((S, n, K), np.float64), np.empty((S, n), np.int32)
        ent, pres = np.empty((S, n), np.float64), np.empty((S, n), np.uint8)
        _check(self._lib.arb_base_cells(self._h, S, K, anc, tm, ll.ctypes.data if want_loglik else None, post, idx,
                                        ent, pres))
        return ll, post, idx, ent, pres.astype(bool)

    def ingest_csr(self, cell_ptr, snv_idx, log_likes, n_cells, n_snvs):
        snv_idx = _i32(snv_idx)
        _check(self._lib.arb_ingest_csr(self._h, int(n_cells), int(n_snvs), snv_idx.shape[0],
                                        _i32(cell_ptr), snv_idx, _f64(log_likes)))

    def dims(self):
        n, m, nnz, npairs = C.c_int32(), C.c_int32(), C.c_int64(), C.c_int32()
        _check(self._lib.arb_get_dims(self._h, C.byref(n), C.byref(m), C.byref(nnz), C.byref(npairs)))
        return n.value, m.value, nnz.value, npairs.value

    def orders(self):
        n, m, _, _ = self.dims()
        co, so = np.empty(n, np.int32), np.empty(m, np.int32)
        _check(self._lib.arb_get_orders(self._h, co, so))
        return co, so

    def count_pairs(self):
        k = self.dims()[3]
        a, t = np.empty(k, np.int32), np.empty(k, np.int32)
        _check(self._lib.arb_get_count_pairs(self._h, a, t))
        return a, t

    def loglik_table(self):
        k = self.dims()[3]
        lp, la = np.empty(k), np.empty(k)
        _check(self._lib.arb_get_loglik_table(self._h, lp, la))
        return lp, la

    def set_loglik_table(self, lp, la):
        lp, la = _f64(lp), _f64(la)
        _check(self._lib.arb_set_loglik_table(self._h, lp.shape[0], lp, la))

    def export_sparse(self, snv_major=False):
        n, m, nnz, _ = self.dims()
        ptr = np.empty((m if snv_major else n) + 1, np.int32)
        idx, ll = np.empty(nnz, np.int32), np.empty(2 * nnz)
        _check(self._lib.arb_export_sparse(self._h, int(bool(snv_major)), ptr, idx, ll))
        return ptr, idx, ll

    # ---- prior ------------------------------------------------------------------
    def set_prior_labels(self, label_idx, n_clusters, gamma):
        _check(self._lib.arb_set_prior_labels(self._h, _i32(label_idx), int(n_clusters), float(gamma)))

    def set_prior(self, q_y_init):
        q = _f64(q_y_init)
        _check(self._lib.arb_set_prior(self._h, q, q.shape[1]))

    def prior(self, n_clusters):
        q = np.empty((self.dims()[1], n_clusters))
        _check(self._lib.arb_get_prior(self._h, q))
        return q

    # ---- fit ----------------------------------------------------------------------
    def posterior_bytes(self, n_trees, n_clones):
        b, f = C.c_int64(), C.c_int64()
        _check(self._lib.arb_posterior_bytes(self._h, int(n_trees), int(n_clones), C.byref(b), C.byref(f)))
        return b.value, f.value

    def fit(self, parents=None, presence=None, n_clones=None, normal_idx=0, max_iter=10,
            tolerance=1.0, keep_posteriors="auto", mode="vi", trace=False):
        """Fit a shard of trees; returns ``(elbo, iters)`` (+ ``trace`` if asked)."""
        if (parents is None) == (presence is None):
            raise ValueError("give exactly one of parents / presence")
        K = int(n_clones)
        T = int(parents.shape[0] if parents is not None else presence.shape[0])
        keep = self._keep_flag(keep_posteriors, T, K)
        elbo = np.empty(T)
        iters = np.empty(T, np.int32)
        tr = np.empty((T, int(max_iter) + 1)) if trace else None
        tr_ptr = tr.ctypes.data if trace else None
        m = {"vi": 0, "ml": 1}[mode]
        self.n_clones = K
        if T == 0:                      # an empty shard (fewer trees than ranks): nothing to launch
            return (elbo, iters, tr) if trace else (elbo, iters)
        if parents is not None:
            _check(self._lib.arb_fit_trees(self._h, T, K, int(normal_idx), _i32(parents), int(max_iter),
                                           float(tolerance), m, keep, elbo, iters, tr_ptr))
        else:
            _check(self._lib.arb_fit_presence(self._h, T, K, np.ascontiguousarray(presence, np.uint8),
                                              int(max_iter), float(tolerance), m, keep, elbo, iters,
                                              tr_ptr))
        self.n_clones = K
        return (elbo, iters, tr) if trace else (elbo, iters)

    def _keep_flag(self, keep_posteriors, T, K):
        if isinstance(keep_posteriors, str):
            if keep_posteriors == "all":
                return 1
            if keep_posteriors == "none":
                return 0
            need, free = self.posterior_bytes(T, K)
            return int(need < 0.35 * free)
        return int(bool(keep_posteriors))

    def get_fit(self, t):
        n, m, _, _ = self.dims()
        qz, qy = np.empty((n, self.n_clones)), np.empty((m, self.n_clones - 1))
        _check(self._lib.arb_get_fit(self._h, int(t), qz, qy))
        return qz, qy

    def get_fits(self, t0, count, fetch=True):
        """Posteriors of trees ``t0 .. t0+count-1`` of the last fit as ``([count, n, K], [count, m, K-1])``.  When
        they were not kept on the device the range is re-fitted as ONE batch and stays cached there
        (``fetch=False`` only does that: later ``get_fit`` / ``get_map`` calls inside the range are copies)."""
        if not fetch:
            _check(self._lib.arb_get_fits(self._h, int(t0), int(count), None, None))
            return None
        n, m, _, _ = self.dims()
        qz = np.empty((count, n, self.n_clones))
        qy = np.empty((count, m, self.n_clones - 1))
        _check(self._lib.arb_get_fits(self._h, int(t0), int(count), qz.ctypes.data, qy.ctypes.data))
        return qz, qy

    def get_map(self, t):
        n, m, _, _ = self.dims()
        z, y = np.empty(n, np.int32), np.empty(m, np.int32)
        _check(self._lib.arb_get_map(self._h, int(t), z, y))
        return z, y

    def get_presence(self, t):
        p = np.empty((self.n_clones - 1, self.n_clones), np.uint8)
        _check(self._lib.arb_get_presence(self._h, int(t), p))
        return p

    # ---- kernel-level hooks ----------------------------------------------------------
    def qz_update(self, presence, q_y):
        P = np.ascontiguousarray(np.asarray(presence) != 0, np.uint8)
        K = P.shape[1]
        qz = np.empty((self.dims()[0], K))
        _check(self._lib.arb_qz_update(self._h, K, P, _f64(q_y), qz))
        return qz

    def qy_update(self, presence, q_z):
        P = np.ascontiguousarray(np.asarray(presence) != 0, np.uint8)
        K = P.shape[1]
        qy = np.empty((self.dims()[1], K - 1))
        _check(self._lib.arb_qy_update(self._h, K, P, _f64(q_z), qy))
        return qy

    def elbo_terms(self, presence, q_z, q_y):
        P = np.ascontiguousarray(np.asarray(presence) != 0, np.uint8)
        lik, ent, kl = C.c_double(), C.c_double(), C.c_double()
        _check(self._lib.arb_elbo_terms(self._h, P.shape[1], P, _f64(q_z), _f64(q_y), C.byref(lik),
                                        C.byref(ent), C.byref(kl)))
        return lik.value, ent.value, kl.value

    # ---- counters / knobs ---------------------------------------------------------------
    def last_fit_stats(self):
        n, ms, ti, sw = C.c_int64(), C.c_double(), C.c_double(), C.c_double()
        _check(self._lib.arb_last_fit_stats(self._h, C.byref(n), C.byref(ms), C.byref(ti), C.byref(sw)))
        return dict(kernel_launches=n.value, device_ms=ms.value, tree_iterations=ti.value,
                    sweep_ms=sw.value)

    def last_fit_profile(self):
        v = np.zeros(24)
        _check(self._lib.arb_last_fit_profile(self._h, v, 24))
        keys = ("sweep_z_ms", "sweep_y_ms", "finish_ms", "tree_sweeps_z", "tree_sweeps_y",
                "launches_z", "launches_y", "rounds", "finish_z_ms", "finish_y_ms", "kp", "panel_rows",
                "panels_z", "panels_y", "tiles_per_warp_z", "tiles_per_warp_y", "fold_z", "fold_y",
                "quads_z", "quads_y", "idle_rounds", "w16", "host_waits", "refit_batches")
        return dict(zip(keys, v.tolist()))

    def set_option(self, name, value):
        _check(self._lib.arb_set_option(self._h, name.encode(), int(value)))


def fit_multi(engines, parents, n_clones, normal_idx=0, max_iter=10, tolerance=1.0, keep_posteriors="auto",
              mode="vi", trace=False):
    """``arb_fit_trees_multi``: one fit spread over several engines (one per GPU, all holding the same matrix and
    prior).  Tree ``t`` is fitted by ``engines[t % G]`` as its local tree ``t // G``; returns ``(elbo, iters)``
    (+ ``trace``) in global tree order, bit-identical to a single-engine fit."""
    lib = load_library()
    G, K = len(engines), int(n_clones)
    parents = _i32(parents)
    T = int(parents.shape[0])
    per = (T + G - 1) // G
    keep = min(e._keep_flag(keep_posteriors, per, K) for e in engines)
    elbo, iters = np.empty(T), np.empty(T, np.int32)
    tr = np.empty((T, int(max_iter) + 1)) if trace else None
    for e in engines:
        e.n_clones = K
    if T:
        handles = (_ctx * G)(*[e._h for e in engines])
        _check(lib.arb_fit_trees_multi(handles, G, T, K, int(normal_idx), parents, int(max_iter), float(tolerance),
                                       {"vi": 0, "ml": 1}[mode], keep, elbo, iters,
                                       tr.ctypes.data if trace else None))
    return (elbo, iters, tr) if trace else (elbo, iters)
