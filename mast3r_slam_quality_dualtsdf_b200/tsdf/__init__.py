"""Mirror of the reference package `mast3r_slam.tsdf` (tsdf/__init__.py:3)."""
from .global_volume import TSDFVolume  # noqa: F401
