"""mola_b200 — B200-native voxel-hashed metric point maps (the hot path of mola_metric_maps).

Host classes mirror the reference's operator interface (names, argument meaning, error
behaviour) on top of the C ABI in include/mola_b200/mmg.h; all compute is hand-written
sm_100a CUDA in mola_b200/csrc. There is no CPU fallback.
"""
from .maps import (HashedVoxelPointCloud, SparseVoxelPointCloud, SparseTreesPointCloud,  # noqa: F401
                   MapError, pose_from_xyz_ypr)

__all__ = ["HashedVoxelPointCloud", "SparseVoxelPointCloud", "SparseTreesPointCloud", "MapError",
           "pose_from_xyz_ypr"]
