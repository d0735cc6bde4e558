"""``arborist.arborist_base`` module path: the reference's legacy scorer
(/root/reference/src/arborist/arborist_base.py) served by the B200 implementation."""
from arborist_b200.arborist_base import (  # noqa: F401
    build_genotypes, find_cell_assignments, main, parse_arguments, rank_trees, save_genotypes, save_tree)

if __name__ == "__main__":
    main()
