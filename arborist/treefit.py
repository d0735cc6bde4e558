"""``arborist.treefit.TreeFit`` -- the path reference pickles name."""
from arborist_b200.treefit import TreeFit  # noqa: F401
