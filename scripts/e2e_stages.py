"""Stage timings of arborist(...) on the headline workload (host wall clock, one GPU)."""
import cProfile, pstats, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from arborist_b200 import arborist, synth

name = sys.argv[1] if len(sys.argv) > 1 else "config3"
d = synth.generate(synth.CONFIGS[name])
rc, cl = d.read_counts_frame(), d.clusters_frame()
kw = dict(alpha=0.001, max_iter=25, tolerance=1.0, gamma=0.7, add_normal=True, device=0)
for _ in range(2):
    lik, best, fits = arborist(d.trees, rc, cl, **kw)
    _ = best.q_z
    arborist.last_stats = None
    del fits, best
t0 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
lik, best, fits = arborist(d.trees, rc, cl, **kw)
qz = best.q_z
pr.disable()
print("wall ms", 1e3 * (time.perf_counter() - t0))
print("device ms", arborist.last_stats["engine"]().last_fit_stats()["device_ms"])
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
