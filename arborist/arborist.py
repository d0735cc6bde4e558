"""``arborist.arborist`` module path (console script target ``arborist.arborist:main``)."""
from arborist_b200.arborist import *  # noqa: F401,F403
from arborist_b200.arborist import arborist, main, parse_arguments, run_variational_inference  # noqa: F401

if __name__ == "__main__":
    main()
