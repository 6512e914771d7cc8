"""Mirror of the reference package `mast3r_slam.tsdf` (tsdf/__init__.py:3)."""
from .global_manager import TSDFGlobalManager  # noqa: F401
from .global_volume import TSDFVolume  # noqa: F401
from .tsdf_optimizer import TSDFPoseOptimizer  # noqa: F401
from .tsdf_factor import TSDFFactor  # noqa: F401  (roadmap interface, MASt3R_SLAM_TSDF_Integration.md 9.3.2)
