"""Drop-in module name for the reference's `scene/ptgs/shen_partation_densify.py`:
put this directory on PYTHONPATH in place of `scene/ptgs` and
`import shen_partation_densify as shen_partition_new` (shen_partition_utils.py:7)
resolves to the B200 implementation.  Pickles written through this module carry
the class path `shen_partation_densify.Partition`, like the reference's
(SURVEY.md A11)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ptgs_b200  # noqa: E402,F401
from ptgs_b200.partition import (  # noqa: E402,F401
    BasicPointCloud, Partition, CameraPose, SimpleCamera, Rect, Point, box,
    quad_tree_partition, expand_partitions, assign_cameras_to_partitions, downsample_point_cloud,
    visibility_based_camera_selection, save_partition_as_pkl, save_partition_data,
    plot_quad_tree, plot_quad_tree_extend, plot_subregions)

for _t in (BasicPointCloud, Partition, CameraPose):
    _t.__module__ = __name__
