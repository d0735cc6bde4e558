"""Row f1 timing: --pickle of ALL fits of the headline workload (config 3, 1,000 trees, 2.7 GB of posteriors)
through the streaming writer, with the posteriors kept on the device and with keep_posteriors="none" (batched
re-fits, what config 4 always needs).  gpurun -- 'python scripts/writers_timing.py [config] [trees]'"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arborist_b200 import arborist, synth

name = sys.argv[1] if len(sys.argv) > 1 else "config3"
d = synth.generate(synth.CONFIGS[name])
trees = d.trees[: int(sys.argv[2])] if len(sys.argv) > 2 else d.trees
rc, cl = d.read_counts_frame(), d.clusters_frame()
out = "/dev/shm/arb_fits.pkl"
for keep in ("all", "none"):
    t0 = time.perf_counter()
    lik, best, fits = arborist(trees, rc, cl, max_iter=25, add_normal=True, device=0, keep_posteriors=keep)
    t1 = time.perf_counter()
    fits.to_pickle(out)
    t2 = time.perf_counter()
    store, eng = arborist.last_stats["store"](), arborist.last_stats["engine"]()
    print(f"{name}, {len(trees)} trees, keep_posteriors={keep}: arborist() {t1 - t0:.2f} s, streaming pickle {t2 - t1:.2f} s "
          f"({os.path.getsize(out) / 1e9:.2f} GB, {os.path.getsize(out) / 1e6 / (t2 - t1):.0f} MB/s), "
          f"batches made resident {store.fetches} (batch = {store.batch} trees), re-fit batches "
          f"{int(eng.last_fit_profile()['refit_batches'])}")
    os.remove(out)
    del fits, best, store, eng
