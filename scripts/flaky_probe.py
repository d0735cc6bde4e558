"""Debugging aid: repeat the K=3 small-workload fit after a fit of another shape in the same process and
compare the ELBOs with the CPU oracle (run under ARB_POOL=0 / ARB_POOL_POISON=1 to localise stale-memory reads)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from arborist_b200 import arborist, synth
from oracle import oracle as orc

def case(k, seed):
    cfg = synth.SynthConfig(f"k{k}", 150, 400, k, 5, seed=seed, coverage=0.08, min_proportion=0.01)
    d = synth.generate(cfg)
    return d, d.read_counts_frame(), d.clusters_frame()

d0, rc0, cl0 = case(10, 7010)
d, rc, cl = case(3, 7003)
olik, obest, ofits, _ = orc.arborist_oracle(d.trees, rc, cl, max_iter=6, tolerance=1.0, add_normal=True)
bad = 0
oe = np.array([olik[t] for t in sorted(olik)])
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    if rep % 2 == 0:
        arborist(d0.trees, rc0, cl0, max_iter=6, tolerance=1.0, add_normal=True)
    lik, best, fits = arborist(d.trees, rc, cl, max_iter=6, tolerance=1.0, add_normal=True)
    e = np.array([lik[t] for t in sorted(lik)])
    ok = best.tree_idx == obest
    bad += not ok
    print(rep, "best", best.tree_idx, obest, "iters", list(arborist.last_stats["iters"]), "oiters",
          [ofits[t][3] for t in sorted(ofits)], "relerr", (np.abs(e - oe) / np.abs(oe)).tolist(), flush=True)
print("bad", bad)
