"""pytest configuration: the ``gpu`` marker and shared golden-fixture helpers."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, f"{name}.npz"), allow_pickle=True)


def golden_frames(g):
    """(trees, read_counts, snv_clusters) as given to ``arborist()`` when the fixture stores them."""
    import pandas as pd

    trees = [[(int(u), int(v)) for u, v in t] for t in g["trees"]]
    rc = pd.DataFrame({"snv": g["rc_snv"], "cell": g["rc_cell"], "alt": g["rc_alt"],
                       "total": g["rc_total"]})
    cl = pd.DataFrame({"snv": g["cl_snv"], "cluster": g["cl_label"]})
    return trees, rc, cl


_LAST_SYNTH = {}


def synth_frames(name, tree_subset=None):
    """Inputs regenerated from the seeded generator (fixtures too big to store verbatim)."""
    from arborist_b200 import synth

    if name not in _LAST_SYNTH:                 # one data set is kept: neighbouring tests at config 4 (a minute of host
        _LAST_SYNTH.clear()                     # time and 3 GB to generate) share it, nothing else accumulates
        _LAST_SYNTH[name] = synth.generate(synth.CONFIGS[name])
    d = _LAST_SYNTH[name]
    trees = d.trees if tree_subset is None else [d.trees[i] for i in tree_subset]
    return trees, d.read_counts_frame(), d.clusters_frame()


@pytest.fixture(scope="session")
def has_cuda():
    try:
        from arborist_b200.engine import device_count
        return device_count() > 0
    except Exception:
        return False
