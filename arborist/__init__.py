"""Drop-in ``arborist`` package name.

Re-exports the B200 implementation under the reference's import paths
(``/root/reference/src/arborist/__init__.py:1-2``): ``arborist.read_trees``,
``arborist.arborist`` (the function), ``arborist.treefit.TreeFit`` (the class
pickles refer to) and the console entry point ``arborist.arborist:main``.
"""
from arborist_b200.utils import read_trees  # noqa: F401
from arborist_b200.arborist import arborist  # noqa: F401,E402  (rebinds the submodule name, as upstream does)
