"""arborist_b200 -- B200-native per-tree variational fit for Arborist.

Public surface mirrors the reference package (``read_trees``, ``arborist``);
the CUDA library behind it is loaded on first use and its absence is an error,
never a CPU fallback.
"""
from .utils import read_trees  # noqa: F401
from .treefit import TreeFit  # noqa: F401
from .arborist import arborist, run_variational_inference, run_simple_max_likelihood  # noqa: F401

__version__ = "0.1.0"
