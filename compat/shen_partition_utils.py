"""Drop-in for `scene/ptgs/shen_partition_utils.py`:
    python compat/shen_partition_utils.py <sfm_path> [threshold_value]
reads <sfm>/sparse/0, writes <sfm>/model/split_result/visible/<id>/... and
partition_<id>.pkl exactly where the reference does (README.md:116-123)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import shen_partation_densify  # noqa: E402,F401  (sets the pickle class paths)
from ptgs_b200.partition import ProgressiveDataPartitioning, main, read_scene  # noqa: E402,F401

if __name__ == "__main__":
    main()
