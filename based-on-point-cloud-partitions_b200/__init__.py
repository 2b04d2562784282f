"""B200-native point-count scene partitioner (drop-in for the hot path of
92401/Based-on-point-cloud-partitions: scene/ptgs/shen_partition_utils.py and
the functions it calls).  Import as `ptgs_b200` through the shim at the repo
root (the directory name is not a Python identifier)."""
__version__ = "0.1.0"
