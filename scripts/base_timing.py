"""Legacy scorer (arborist_base.rank_trees) on config 3: wall time on the GPU and the numpy oracle beside it."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arborist_b200 import arborist_base as ab, synth
from oracle import oracle_base as ob

name = sys.argv[1] if len(sys.argv) > 1 else "config3"
trees, rc = synth.legacy_inputs(name)
for rep in range(3):
    t0 = time.perf_counter()
    ranked, cells = ab.rank_trees(trees, rc, alpha=0.001, topn=25)
    print(f"gpu rank_trees {len(trees)} trees, {len(rc)} rows: {1e3 * (time.perf_counter() - t0):.1f} ms", flush=True)
t0 = time.perf_counter()
oranked, _ = ob.rank_trees_oracle(trees, rc, 0.001, 25)
print(f"numpy oracle: {1e3 * (time.perf_counter() - t0):.1f} ms; same ranking: {ranked['tree'].tolist() == oranked['tree'].tolist()}; "
      f"max rel diff {float((abs(ranked.sort_values('tree')['likelihood'].to_numpy() - oranked.sort_values('tree')['likelihood'].to_numpy()) / abs(oranked.sort_values('tree')['likelihood'].to_numpy())).max()):.2e}")
