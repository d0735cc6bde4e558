"""Paths that need MORE than one physical GPU (skipped on a one-GPU box; run with ``gpurun --gpus 2``):
one process per GPU over NCCL through ``arborist()`` exactly as INTEGRATION.md says -- no
``torch.cuda.set_device`` in the caller -- and one process driving several GPUs (``arb_fit_trees_multi``).
Both must return the bits of a one-GPU run (reference loop A.py:1089-1126: trees are independent)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import ROOT, synth_frames


def _gpus():
    from arborist_b200.engine import device_count
    try:
        return device_count()
    except Exception:
        return 0


NCCL_WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, os.environ["ARB_ROOT"]); sys.path.insert(0, os.path.join(os.environ["ARB_ROOT"], "tests"))
    from conftest import synth_frames
    from arborist_b200 import arborist
    dist.init_process_group("nccl")                  # and NO torch.cuda.set_device: arborist() has to find its GPU
    rank = dist.get_rank()
    trees, rc, cl = synth_frames("small")
    lik, best, fits = arborist(trees, rc, cl, add_normal=True)
    assert arborist.last_stats["devices"] == [int(os.environ["LOCAL_RANK"])]
    np.save(os.path.join(os.environ["ARB_OUT"], "elbo%d.npy" % rank), np.array([lik[i] for i in range(len(trees))]))
    # a tree only rank 1 validates is broken (cycle -> StopIteration): every rank raises, nobody hangs
    bad = [trees[0], [(1, 2), (2, 1)]] + trees[2:4]
    try:
        arborist(bad, rc, cl, add_normal=True)
        raise SystemExit("no error on rank %d" % rank)
    except (ValueError, StopIteration):
        pass
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(os.environ["ARB_OUT"], "rank%d.ok" % rank), "w").write("ok")
""")


@pytest.mark.gpu
def test_nccl_two_ranks_without_set_device(tmp_path):
    if _gpus() < 2:
        pytest.skip("needs two GPUs")
    from arborist_b200 import arborist
    trees, rc, cl = synth_frames("small")
    lik, _, _ = arborist(trees, rc, cl, add_normal=True, device=0)
    want = np.array([lik[i] for i in range(len(trees))])
    script = tmp_path / "worker.py"
    script.write_text(NCCL_WORKER)
    env = dict(os.environ, ARB_ROOT=ROOT, ARB_OUT=str(tmp_path))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    for r in (0, 1):
        assert (tmp_path / f"rank{r}.ok").exists()
        assert (np.load(tmp_path / f"elbo{r}.npy") == want).all()          # every rank holds the full, identical vector


@pytest.mark.gpu
def test_one_process_on_every_gpu_gives_the_one_gpu_bits():
    n = _gpus()
    if n < 2:
        pytest.skip("needs two GPUs")
    from arborist_b200 import arborist
    trees, rc, cl = synth_frames("config2")
    lik1, best1, _ = arborist(trees, rc, cl, add_normal=True, device=0)
    likn, bestn, fitsn = arborist(trees, rc, cl, add_normal=True, devices="all")
    assert arborist.last_stats["devices"] == list(range(n))
    assert [lik1[i] for i in range(len(trees))] == [likn[i] for i in range(len(trees))]
    assert best1.tree_idx == bestn.tree_idx
    assert (best1.q_z == bestn.q_z).all() and (best1.q_y == bestn.q_y).all()
    last = fitsn[len(trees) - 1]                       # a tree another GPU fitted
    assert np.isfinite(last.q_z).all()
