"""Tree-set sharding over ranks (world_size 2, gloo, CPU): shard indices, ELBO gather and the
owner-broadcast of posteriors, with a stand-in engine (the GPU engine is covered by -m gpu)."""
import os
import subprocess
import sys
import textwrap

import numpy as np

from conftest import ROOT

WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, os.environ["ARB_ROOT"])
    from arborist_b200 import distributed as D

    dist.init_process_group("gloo")
    rank, ws = D.world()
    assert ws == 2
    T = 7
    shard = D.shard_indices(T)
    assert shard.tolist() == list(range(rank, T, 2))
    truth = np.arange(T) * 1.5 - 3.0
    full, owner = D.gather_elbos(truth[shard], shard, T)
    assert np.array_equal(full, truth), full
    assert owner.tolist() == [t % 2 for t in range(T)]
    iters = D.gather_vector((shard + 3).astype(np.int32), shard, T, dtype=np.int32)
    assert iters.tolist() == [t + 3 for t in range(T)]

    from arborist_b200.results import FitStore, LazyFits, write_fits_pickle
    from arborist_b200.treefit import TreeFit

    class FakeEngine:
        n_clones, device = 3, 0
        def dims(self): return 4, 5, 0, 0
        def get_fits(self, t0, count, fetch=True):
            assert not fetch and t0 % 2 == 0 and count <= 2          # batches of max_batch consecutive local trees
        def get_fit(self, t):
            g = float(shard[t])
            return np.full((4, 3), g), np.full((5, 2), -g)

    store = FitStore([FakeEngine()], shard, owner, T, max_batch=2)
    for idx in (0, 1, 6):           # collective: both ranks ask for the same tree
        qz, qy = store.get_fit(idx)
        assert (qz == idx).all() and (qy == -idx).all() and qz.shape == (4, 3)
    assert D.all_agree(True) and not D.all_agree(rank == 0)
    # arborist(): a tree that only rank 1's shard validates is broken in a way that is NOT a ValueError (a cycle has
    # no root: StopIteration).  Rank 0 must not be left alone in the collective: both ranks raise, before any engine.
    import pandas as pd
    from arborist_b200 import arborist
    rc = pd.DataFrame({"snv": [1, 2], "cell": [1, 1], "alt": [1, 0], "total": [1, 1]})
    cl = pd.DataFrame({"snv": [1, 2], "cluster": [1, 2]})
    try:
        arborist([[(1, 2)], [(1, 2), (2, 1)], [(1, 2)]], rc, cl, add_normal=True)
        raised = None
    except Exception as e:
        raised = type(e).__name__
    assert raised == ("ValueError" if rank == 0 else "StopIteration"), raised
    assert D.local_device(3) == 3 and D.local_device() == 0          # no GPU here: LOCAL_RANK % max(1, 0)
    # streaming pickle of all fits: every rank takes part (remote posteriors are broadcast), rank 0 writes
    dicts = [{"a": 0, "b": 1, "c": 2, "d": 3}, {i: i for i in range(5)}, {0: 0, 1: 1, 2: 2}, {1: 0, 2: 1}]
    fits = LazyFits(T, lambda i: TreeFit.lazy([(0, 1), (1, 2)], i, float(truth[i]), store.get_fit, *dicts), store)
    path = os.path.join(os.environ["ARB_OUT"], "fits.pkl")
    write_fits_pickle(fits, path)
    dist.barrier()
    if rank == 0:
        import pandas as pd
        back = pd.read_pickle(path)
        assert sorted(back) == list(range(T))
        assert all((back[i].q_z == i).all() and back[i].elbo == truth[i] for i in range(T))
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(os.environ["ARB_OUT"], "rank%d.ok" % rank), "w").write("ok")
""")


def test_gloo_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, ARB_ROOT=ROOT, ARB_OUT=str(tmp_path), OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert (tmp_path / "rank0.ok").exists() and (tmp_path / "rank1.ok").exists()


def test_single_process_defaults():
    from arborist_b200 import distributed as D
    assert D.world() == (0, 1)
    assert D.shard_indices(5).tolist() == [0, 1, 2, 3, 4]
    full, owner = D.gather_elbos(np.array([1.0, 2.0]), np.array([0, 1]), 2)
    assert full.tolist() == [1.0, 2.0] and owner.tolist() == [0, 0]
