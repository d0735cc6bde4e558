#!/usr/bin/env python
"""bench.py -- candidate trees fit+ranked per second on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over the whole candidate set of the headline workload
(BASELINE.json configs[2]: synthetic 10k cells x 20k SNVs low-pass, 12 clones, 1,000 candidate
trees per GPU): batched CAVI fit of every tree (q(y)/q(z) sweeps + ELBO), the NCCL gather of
the ELBOs when N > 1, and the ranking.  Trees are independent, so the tree set is sharded over
ranks with no data-path collective; per-GPU work is fixed as N grows ("weak").

* ``value``    trees/s with the read-count matrix already resident in HBM (device events).
* ``e2e``      the same metric through the public ``arborist(...)`` call with HOST DataFrames:
               id factorisation, H2D of the COO table, device ingest, fit, D2H of ELBOs and
               of the best tree's posteriors are all inside the timed region.
* ``roofline`` q(z) sweep kernel: algorithmic bytes / CUDA-event duration,
               against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
* ``cpu_baseline`` the CPU oracle (C port of the reference's numba kernels, OpenMP on all
               host cores) on a bounded sample of the same workload.

``--impl reference`` times that CPU implementation alone (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "candidate_trees_fit_and_ranked_per_second"
RUN = dict(alpha=0.001, gamma=0.7, max_iter=25, tolerance=1.0)      # CLI defaults, A.py:62-81
# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per tree-sweep of the two sweep kernels, one entry per
# config that has an `ncu --set full` capture under profiles/ (per launch the kernel moves this times the
# tree-sweeps of the launch; the figure is NOT measured in the bench run itself -- ncu cannot run inside a timed
# region).  Configs without a capture report traffic = null with the reason.
NCU_DRAM_BYTES_PER_TREE_SWEEP = {
    "config3": {"z": 28.2e6, "y": 14.1e6, "capture": "profiles/r2_full_summary.md (r2b)"},
    "config5": {"z": 17.2e6, "y": 8.0e6, "capture": "profiles/r2_full_summary.md (r2k5)"},
    "config4": {"z": 913e6, "y": 471e6, "capture": "profiles/r2_full_summary.md (r2k4)"},
    # 100 trees in flight, second round: 29.83 MB per q(z) sweep launch, 6.66 MB per q(y) sweep launch -- the matrix,
    # the operands and the partial rows of all 100 trees stay in the 126 MB L2, DRAM sees little of the 2 MB a
    # tree-sweep moves algorithmically
    "config2": {"z": 0.298e6, "y": 0.0666e6, "capture": "profiles/r3d_config2_summary.md"},
}


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--config", default="config3")
    p.add_argument("--trees", type=int, default=0, help="candidate trees per GPU (default: the config's)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--slots", type=int, default=0)
    p.add_argument("--lookahead", type=int, default=0, help="rounds the host enqueues ahead (0 = library default)")
    p.add_argument("--split", type=int, default=0, help="force the sweep split (0 = chosen on the device)")
    p.add_argument("--no-single", action="store_true", help="skip the single-process leg (N > 1: after the timed legs "
                   "rank 0 alone drives every GPU through arborist(..., devices='all') = arb_fit_trees_multi)")
    p.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg (N > 1: the config's own "
                   "candidate set, e.g. 1,000 trees in total, sharded over the N GPUs)")
    return p.parse_args()


# ---------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------
def workload_name(cfg, n, m, nnz, K, per, total):
    return (f"{cfg.name}: {cfg.n_cells} cells x {cfg.n_snvs} SNVs low-pass ({n} x {m} after the zero-alt "
            f"filter, nnz {nnz}), K={K} clones, {per} candidate trees per GPU ({total} total), seed {cfg.seed}")


def make_workload(name, trees_per_gpu, n_gpus):
    """The seeded synthetic dataset + a global candidate set of ``trees_per_gpu * n_gpus`` trees
    (the config's own trees first, then further seeded random trees over the same labels)."""
    from arborist_b200 import synth

    cfg = synth.CONFIGS[name]
    d = synth.generate(cfg)
    per = trees_per_gpu or cfg.n_trees
    want = per * n_gpus
    trees = list(d.trees[:want])
    rng = np.random.default_rng(cfg.seed + 104729)
    while len(trees) < want:
        trees.append(synth._edges(synth.random_tree(rng, cfg.n_clones - 1, cfg.deep)))
    return cfg, d, trees, per


def algorithmic_bytes(n, m, nnz, K):
    """bytes one tree-iteration must move at the reference's element widths."""
    C = K - 1
    b_z = 20 * nnz + 4 * (n + 1) + 8 * (n * K + m * C)            # CSR sweep: q(z) update + ELBO
    b_y = 20 * nnz + 4 * (m + 1) + 8 * (n * K + 2 * m * C)        # CSC sweep: q(y) update
    return b_z, b_y


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle (C port of the reference's kernels) on the host cores
# ---------------------------------------------------------------------------------------------
class CpuArm:
    """One WHOLE candidate tree per step, fitted to convergence as the reference's loop does (A.py:1089-1126 ->
    run_variational_inference, A.py:679-816, with RUN's max_iter / tolerance) by the C/OpenMP oracle on every thread
    the host offers.  The reference's ingest (A.py:1034-1069 = ``orc.prepare``) runs once and is timed apart."""

    def __init__(self, cfg, d, trees):
        from oracle import oracle as orc

        self.orc, self.trees = orc, trees
        # every core this process may run on (torchrun exports OMP_NUM_THREADS=1: do not inherit that)
        try:
            self.cores = len(os.sched_getaffinity(0))
        except AttributeError:
            self.cores = os.cpu_count() or 1
        orc.lib().orc_set_threads(self.cores)
        t0 = time.perf_counter()
        self.prob = orc.prepare(trees[:1], d.read_counts_frame(), d.clusters_frame(), alpha=RUN["alpha"],
                                gamma=RUN["gamma"], add_normal=True)
        self.prepare_s = time.perf_counter() - t0
        self.set_size = cfg.n_trees
        self.add_normal = orc.add_normal_clone
        self.iters = []
        self.sample = (f"one whole candidate tree per step (tree number warmup + s of the set), fitted to convergence "
                       f"(max_iter {RUN['max_iter']}, tolerance {RUN['tolerance']}) by the C/OpenMP port of the "
                       f"reference's kernels on {self.cores} threads; trees / seconds, nothing extrapolated")

    def step(self, s):
        tree = self.add_normal([self.trees[s % len(self.trees)]])[0]
        t0 = time.perf_counter()
        _, _, _, it, _ = self.orc.fit_tree(self.prob, tree, max_iter=RUN["max_iter"], tolerance=RUN["tolerance"],
                                           return_trace=True)
        dt = time.perf_counter() - t0
        self.iters.append(int(it))
        return dt, it + 0.5


def bench_config(cfg, n, m, nnz, K, per, total, world):
    """The ``config`` object of the JSON line -- one function so that both arms print the same object."""
    return {"workload": workload_name(cfg, n, m, nnz, K, per, total),
            "run": RUN, "parallelism": f"tree-sharded x{world}",
            "l2": "inputs larger than L2: per-step working set (panel partials + posteriors of "
                  "all trees in flight) is tens of GB, far above the 126 MB L2; no flush needed",
            "timing": "CUDA events on the library's launching stream around each fit + torch "
                      "events around the ELBO gather; max over ranks"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(1, args.gpus)
    cfg, d, trees, per = make_workload(args.config, args.trees, world)
    arm = CpuArm(cfg, d, trees)
    for s in range(args.warmup):
        arm.step(s)
    arm.iters = []
    tot_t = tot_i = 0.0
    for s in range(args.steps):
        dt, ti = arm.step(args.warmup + s)
        tot_t += dt
        tot_i += ti
    steps = max(1, args.steps)
    v = steps / tot_t
    # end to end = the fit plus the reference's ingest, which one arborist(...) call pays once for its candidate set
    v_e2e = steps / (tot_t + steps * arm.prepare_s / arm.set_size)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "trees/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(cfg, arm.prob.n_cells, arm.prob.n_snvs, len(arm.prob.snv_idx_cs), arm.prob.n_clones,
                               per, per * world, world),
        "step": arm.sample, "iters_per_tree": arm.iters, "tree_iters_per_s": tot_i / tot_t,
        "ingest_s": arm.prepare_s,
        "cpu_baseline": {"value": v, "unit": "trees/s", "cores": arm.cores, "kind": "port", "sample": arm.sample},
        "e2e": {"value": v_e2e, "unit": "trees/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": f"fit time plus the ingest (A.py:1034-1069, {arm.prepare_s:.2f} s here) spread over the "
                        f"{arm.set_size} trees of the config's candidate set"},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch

    from arborist_b200 import arborist, distributed, engine
    from arborist_b200.arborist import add_normal_clone, ingest_read_counts, prepare_trees

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if engine.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist = None

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cfg, d, trees, per = make_workload(args.config, args.trees, world)
    rc, cl = d.read_counts_frame(), d.clusters_frame()
    T = len(trees)

    # ---- resident state: ingest once, keep the matrix in HBM ----------------------------------
    trees_n = add_normal_clone(trees)
    clone_to_idx, cluster_to_idx, parents, presence, normal_idx = prepare_trees(trees_n)
    K = len(clone_to_idx)
    eng = engine.Engine(device=local)
    if args.slots:
        eng.set_option("slots", args.slots)
    if args.lookahead:
        eng.set_option("lookahead", args.lookahead)
    if args.split:
        eng.set_option("split", args.split)
    ingest_read_counts(eng, rc, cl, cluster_to_idx, RUN["alpha"], RUN["gamma"])
    n, m, nnz, n_pairs = eng.dims()
    shard = distributed.shard_indices(T)
    my_parents = np.ascontiguousarray(parents[shard])
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def device_step():
        """fit + gather + rank; returns (device ms, stats, profile)."""
        elbo, iters = eng.fit(parents=my_parents, n_clones=K, normal_idx=normal_idx,
                              max_iter=RUN["max_iter"], tolerance=RUN["tolerance"], keep_posteriors="all")
        st, pf = eng.last_fit_stats(), eng.last_fit_profile()
        ev0.record()
        full, _ = distributed.gather_elbos(elbo, shard, T)
        ev1.record()
        torch.cuda.synchronize()
        order = np.argsort(-full, kind="stable")               # the ranking (A.py:1164-1166)
        return st["device_ms"] + ev0.elapsed_time(ev1), st, pf, full, order, iters

    for _ in range(args.warmup):
        device_step()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    dev_ms = wall = 0.0
    launches = 0
    prof = dict(sweep_z_ms=0.0, sweep_y_ms=0.0, finish_ms=0.0, tree_sweeps_z=0.0, tree_sweeps_y=0.0,
                launches_z=0.0, launches_y=0.0, rounds=0.0, finish_z_ms=0.0, finish_y_ms=0.0)
    tree_iters = 0.0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        ms, st, pf, full, order, iters = device_step()
        dev_ms += ms
        launches += st["kernel_launches"]
        tree_iters += st["tree_iterations"]
        for k in prof:
            prof[k] += pf[k]
    barrier()
    wall = time.perf_counter() - t_wall0
    clocks = sampler.stop()
    dev_ms = max_over_ranks(dev_ms)
    wall = max_over_ranks(wall)
    tree_iters_all = tree_iters
    if dist is not None:
        t = torch.tensor([tree_iters], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        tree_iters_all = float(t.item())
    steps = max(1, args.steps)
    value = T * steps / (dev_ms * 1e-3)

    # ---- identity of the answers across N: hashes of the ELBOs / iteration counts of the config's OWN candidate set --
    # (the first cfg.n_trees trees are the same at every N; which GPU fitted a tree must not change a bit of it)
    import hashlib

    def digest(a, dtype):
        return hashlib.sha256(np.ascontiguousarray(a, dtype=dtype).tobytes()).hexdigest()

    T_h = min(T, cfg.n_trees)
    iters_all = distributed.gather_vector(iters, shard, T, dtype=np.int32)          # untimed
    answers = {"trees": int(T_h), "elbo_sha256": digest(full[:T_h], np.float64),
               "iters_sha256": digest(iters_all[:T_h], np.int32),
               "note": "sha256 of the fp64 ELBO bytes / int32 iteration counts of the config's first `trees` candidate "
                       "trees (the same trees at every N): equal hashes = bit-identical answers whichever GPU fitted them"}

    # ---- strong scaling: the config's OWN candidate set (config 3: 1,000 trees in total) sharded over the ranks ---
    strong = None
    if world > 1 and not args.no_strong:
        T_s = min(T, cfg.n_trees)
        shard_s = distributed.shard_indices(T_s)
        par_s = np.ascontiguousarray(parents[shard_s])

        def strong_step():
            elbo, it_s = eng.fit(parents=par_s, n_clones=K, normal_idx=normal_idx, max_iter=RUN["max_iter"],
                                 tolerance=RUN["tolerance"], keep_posteriors="all")
            ms = eng.last_fit_stats()["device_ms"]
            ev0.record()
            full_s, _ = distributed.gather_elbos(elbo, shard_s, T_s)
            ev1.record()
            torch.cuda.synchronize()
            strong_step.iters = it_s
            return ms + ev0.elapsed_time(ev1), full_s

        for _ in range(2):
            strong_step()
        barrier()
        s_ms = 0.0
        for _ in range(args.steps):
            ms, full_s = strong_step()
            s_ms += ms
        barrier()
        s_ms = max_over_ranks(s_ms)
        iters_s = distributed.gather_vector(strong_step.iters, shard_s, T_s, dtype=np.int32)      # untimed
        # the same through arborist(...) with host DataFrames (every rank: prepare its shard's trees, ingest the table,
        # fit, gather): the per-rank ingest does not shrink with N, so this is the Amdahl view of strong scaling
        s_e2e = None
        if not args.no_e2e:
            def strong_e2e():
                lik, best_s, _ = arborist(trees[:T_s], rc, cl, alpha=RUN["alpha"], max_iter=RUN["max_iter"],
                                          tolerance=RUN["tolerance"], gamma=RUN["gamma"], add_normal=True, device=local)
                return best_s.tree_idx
            strong_e2e()
            barrier()
            t_s0 = time.perf_counter()
            for _ in range(args.steps):
                strong_e2e()
            barrier()
            s_e2e = T_s * steps / max_over_ranks(time.perf_counter() - t_s0)
            arborist.last_stats = None
        strong = {"trees_total": int(T_s), "trees_per_gpu": float(T_s) / world, "value": T_s * steps / (s_ms * 1e-3),
                  "e2e_value": s_e2e,
                  "unit": "trees/s", "ms_per_step": s_ms / steps, "best_tree": int(np.argmax(full_s)),
                  "elbo_sha256": digest(full_s, np.float64), "iters_sha256": digest(iters_s, np.int32),
                  "note": "same metric with the TOTAL candidate set fixed (north_star: 1,000 trees, 1 -> 8 GPUs); "
                          "device events, max over ranks; divide by the N = 1 line's value for the speed-up"}

    # ---- roofline of the dominant kernel (q(z) sweep), from this rank's events --------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    sm_count = torch.cuda.get_device_properties(local).multi_processor_count
    b_z, b_y = algorithmic_bytes(n, m, nnz, K)
    lz = max(1.0, prof["launches_z"])
    ach_z = prof["tree_sweeps_z"] * b_z / max(prof["sweep_z_ms"], 1e-9) / 1e6          # GB/s
    ach_y = prof["tree_sweeps_y"] * b_y / max(prof["sweep_y_ms"], 1e-9) / 1e6
    roofline = {
        "kernel": "k_sweep<KP,true> (q(z) sweep + E[log lik] accumulators)", "bound": "hbm",
        "achieved": ach_z, "peak": peak, "unit": "GB/s", "frac": ach_z / peak,
        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
        "traffic": (NCU_DRAM_BYTES_PER_TREE_SWEEP[cfg.name]["z"] * prof["tree_sweeps_z"] / lz
                    if cfg.name in NCU_DRAM_BYTES_PER_TREE_SWEEP else None),
        "traffic_source": (f"ncu --set full, {NCU_DRAM_BYTES_PER_TREE_SWEEP[cfg.name]['capture']}: "
                           f"{NCU_DRAM_BYTES_PER_TREE_SWEEP[cfg.name]['z'] / 1e6:.1f} MB of DRAM traffic per tree-sweep "
                           "(dram__bytes_read.sum + dram__bytes_write.sum of one captured launch / its tree-sweeps) x "
                           "this run's tree-sweeps per launch; a constant from that capture, not measured in this run"
                           if cfg.name in NCU_DRAM_BYTES_PER_TREE_SWEEP else
                           "null: no ncu --set full capture of this config under profiles/"),
        "algorithmic_bytes_per_tree_sweep": b_z, "tree_sweeps_per_launch": prof["tree_sweeps_z"] / lz,
        "avg_launch_ms": prof["sweep_z_ms"] / lz,
        "qy_sweep": {"achieved": ach_y, "frac": ach_y / peak, "algorithmic_bytes_per_tree_sweep": b_y,
                     "avg_launch_ms": prof["sweep_y_ms"] / max(1.0, prof["launches_y"])},
        "whole_step": {"achieved": tree_iters / steps * (b_z + b_y) / (dev_ms / steps) / 1e6,
                       "frac": tree_iters / steps * (b_z + b_y) / (dev_ms / steps) / 1e6 / peak,
                       "note": "tree-iterations x B_iter / step time, per GPU"},
        "kernel_share_of_step": {"sweep_z": prof["sweep_z_ms"] / (dev_ms if world == 1 else max(dev_ms, 1e-9)),
                                 "sweep_y": prof["sweep_y_ms"] / max(dev_ms, 1e-9),
                                 "finish_z": prof["finish_z_ms"] / max(dev_ms, 1e-9),
                                 "finish_y": prof["finish_y_ms"] / max(dev_ms, 1e-9)},
        # what really bounds the sweeps: the L1 / shared-memory data pipe (one 128-byte wavefront per nonzero and
        # 16 packed columns = 4 chunks), not HBM -- so frac above can neither reach nor be compared with 1.0 across
        # configs: at K clones a tree-sweep needs nnz * ceil(cols/4)/4 wavefronts whatever its algorithmic bytes
        "lsu_floor": (lambda wz, wy, hz: {
            "pipe": "l1tex data pipe: 1 wavefront (128 B) per SM per cycle; a nonzero of a virtual slot (16 packed "
                    "columns) is one wavefront",
            "sm_count": sm_count, "sm_mhz_used": hz,
            "z_frac_of_floor": wz / (sm_count * hz * 1e3) / max(prof["sweep_z_ms"], 1e-9),
            "y_frac_of_floor": wy / (sm_count * hz * 1e3) / max(prof["sweep_y_ms"], 1e-9),
            "max_hbm_frac_at_floor": {"z": (b_z / peak / 1e9) / (nnz * ((K + 3) // 4) / 4 / (sm_count * hz * 1e6)),
                                      "y": (b_y / peak / 1e9) / (nnz * ((K + 2) // 4) / 4 / (sm_count * hz * 1e6))},
        })(prof["tree_sweeps_z"] * nnz * ((K + 3) // 4) / 4.0, prof["tree_sweeps_y"] * nnz * ((K + 2) // 4) / 4.0,
           float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)),
        "layout": {k: int(pf[k]) for k in ("kp", "panel_rows", "panels_z", "panels_y", "tiles_per_warp_z",
                                            "tiles_per_warp_y", "fold_z", "fold_y", "quads_z", "quads_y", "w16")},
    }

    # ---- end to end through arborist(...) with host DataFrames ---------------------------------------
    e2e = None
    if not args.no_e2e:
        eng.close()
        del eng
        e2e_dev_ms, e2e_stages = [], []

        def e2e_step():
            lik, best, fits = arborist(trees, rc, cl, alpha=RUN["alpha"], max_iter=RUN["max_iter"],
                                       tolerance=RUN["tolerance"], gamma=RUN["gamma"], add_normal=True,
                                       device=local)
            ranking = sorted(lik.items(), key=lambda kv: -kv[1])
            qz, qy = best.q_z, best.q_y                           # D2H of the winner's posteriors
            nbytes = 12 * len(shard) + 4 * (n + m) + qz.nbytes + qy.nbytes
            assert ranking[0][0] == best.tree_idx
            e2e_dev_ms.append(round(arborist.last_stats["engine"]().last_fit_stats()["device_ms"], 1))
            e2e_stages.append(dict(arborist.last_stats["stages_ms"], after_call=0.0))
            t_after = time.perf_counter()
            arborist.last_stats = None                            # drop the engine (its buffers go to the pool)
            del fits, best
            e2e_stages[-1]["after_call"] = round(1e3 * (time.perf_counter() - t_after), 2)
            return nbytes

        for _ in range(min(1, args.warmup)):                      # one untimed call, like the device arm's warm-up
            e2e_step()
        import gc
        gc.collect()                             # the start-up heap (module objects of torch / pandas) leaves the collector,
        gc.freeze()                              # as in a long-running service; arborist() itself no longer does this
        gc_log, gc_t0 = [], [0.0]

        def gc_cb(phase, info):
            if phase == "start":
                gc_t0[0] = time.perf_counter()
            else:
                dt = 1e3 * (time.perf_counter() - gc_t0[0])
                if dt > 5:
                    gc_log.append((info["generation"], round(dt, 1)))
        gc.callbacks.append(gc_cb)               # pauses above 5 ms are listed
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        step_ms = []
        for _ in range(args.steps):
            t1 = time.perf_counter()
            d2h = e2e_step()
            step_ms.append(round(1e3 * (time.perf_counter() - t1), 1))
        barrier()
        e2e_wall = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": T * steps / e2e_wall, "unit": "trees/s",
               # the four int64 columns of the table (ids are factorised on the device), prior labels, parent
               # vectors, scipy's log-pmf table
               "h2d_bytes_per_step": int(32 * len(rc) + 4 * m + 4 * my_parents.size + 16 * n_pairs),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * e2e_wall / steps, "ms_steps_rank0": step_ms, "gc_pauses_gen_ms": gc_log,
               "fit_device_ms_steps_rank0": e2e_dev_ms[-len(step_ms):],
               "host_stages_ms_steps_rank0": e2e_stages[-len(step_ms):],
               "note": "arborist(trees, read_counts_df, clusters_df): H2D of the long table (int64 columns), device "
                       "id factorisation + ingest + layouts, batched fit, ELBO gather, ranking, D2H of ELBOs / "
                       "iteration counts / index orders / the best tree's q_z and q_y; host wall clock, one untimed "
                       "warm-up call first (like the device arm's warm-up steps); per-call wall times, the fit's device "
                       "time, host stage times and GC pauses above 5 ms are listed; bench.py runs gc.collect() + gc.freeze() once "
                       "after the warm-up call (start-up heap out of the collector, as a long-running service would; "
                       "arborist() leaves the collector alone unless ARBORIST_GC_FREEZE=1)"}

    # ---- CPU baseline (rank 0, N = 1) ------------------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(cfg, d, trees)
        dt = ti = 0.0
        for s in range(2):                       # trees 0 and 1 of the set, whole: ~25 s on 16 cores at config 3
            a_dt, a_ti = arm.step(s)
            dt += a_dt
            ti += a_ti
        cpu = {"value": 2 / dt, "unit": "trees/s", "cores": arm.cores, "kind": "port",
               "sample": arm.sample.replace("tree number warmup + s", "trees 0 and 1"), "iters_per_tree": arm.iters,
               "tree_iters_per_s": ti / dt, "seconds": dt, "ingest_s": arm.prepare_s}

    # ---- N > 1: the same candidate set through ONE process that drives every GPU itself -------------------------
    # (arborist(..., devices="all") -> arb_fit_trees_multi, the console script's default on a multi-GPU box).  Runs
    # after every timed leg: all ranks give their device memory back and leave the process group, then rank 0 alone
    # opens a context per GPU.  The hashes must equal `answers` / `strong`.
    single = None
    if world > 1 and not args.no_single:
        import gc

        try:
            eng.close()
        except NameError:
            pass
        arborist.last_stats = None
        gc.collect()
        engine.trim_pool()
        torch.cuda.empty_cache()
        barrier()
        dist.destroy_process_group()
        dist = None
        if rank == 0:
            try:
                def single_call():
                    lik_sp, best_sp, _ = arborist(trees[:T_h], rc, cl, alpha=RUN["alpha"], max_iter=RUN["max_iter"],
                                                  tolerance=RUN["tolerance"], gamma=RUN["gamma"], add_normal=True,
                                                  devices="all")
                    return lik_sp, best_sp.tree_idx

                single_call()                                            # untimed: one context per GPU is created here
                t_sp = time.perf_counter()
                for _ in range(steps):
                    lik_sp, best_sp = single_call()
                t_sp = time.perf_counter() - t_sp
                single = {"devices": list(arborist.last_stats["devices"]), "trees": int(T_h),
                          "trees_per_s": T_h * steps / t_sp, "ms_per_call": 1e3 * t_sp / steps,
                          "elbo_sha256": digest([lik_sp[i] for i in range(T_h)], np.float64),
                          "iters_sha256": digest(arborist.last_stats["iters"], np.int32), "best_tree": int(best_sp),
                          "stages_ms_last_call": arborist.last_stats["stages_ms"],
                          "note": "one process, one host thread per GPU (arb_fit_trees_multi), host DataFrames in, "
                                  "wall clock of the whole call; run by rank 0 after the other ranks released their GPUs"}
                arborist.last_stats = None
            except Exception as e:                                        # never lose the line over the extra leg
                single = {"error": repr(e)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "trees/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(cfg, n, m, nnz, K, per, T, world),
            "tree_iters_per_s": tree_iters_all / (dev_ms * 1e-3),
            "mean_tree_iters_per_tree": tree_iters / steps / len(shard),
            "wall_ms_per_step": 1e3 * wall / steps,
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "e2e": e2e, "cpu_baseline": cpu, "strong": strong,
            "round_pipeline": {"host_waits": pf["host_waits"], "idle_rounds": pf["idle_rounds"], "rounds": pf["rounds"]},
            "best_tree": int(order[0]), "true_tree": int(d.true_tree_index), "answers": answers,
            "single_process": single,
        }
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    # keep stdout clean for the ONE json line: libraries (NCCL_DEBUG, torchrun banners) print to fd 1
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
