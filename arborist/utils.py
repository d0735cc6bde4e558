"""``arborist.utils`` -- tree file reader and drawing helper."""
from arborist_b200.utils import read_trees, visualize_tree, write_trees  # noqa: F401
