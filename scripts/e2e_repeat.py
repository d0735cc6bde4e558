"""Wall time of repeated arborist(...) calls on the headline workload (one GPU); ARB_POOL_TRACE=1 shows pool misses."""
import gc, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arborist_b200 import arborist, synth

d = synth.generate(synth.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "config3"])
rc, cl = d.read_counts_frame(), d.clusters_frame()
kw = dict(alpha=0.001, max_iter=25, tolerance=1.0, gamma=0.7, add_normal=True, device=0)
for i in range(int(sys.argv[2]) if len(sys.argv) > 2 else 10):
    t0 = time.perf_counter()
    lik, best, fits = arborist(d.trees, rc, cl, **kw)
    qz = best.q_z
    t1 = time.perf_counter()
    dev = arborist.last_stats["engine"]().last_fit_stats()["device_ms"]
    arborist.last_stats = None
    del fits, best, qz, lik
    t2 = time.perf_counter()
    print(f"call {i}: arborist {1e3*(t1-t0):.1f} ms (device {dev:.1f}), drop {1e3*(t2-t1):.1f} ms, gc counts {gc.get_count()}",
          file=sys.stderr, flush=True)
