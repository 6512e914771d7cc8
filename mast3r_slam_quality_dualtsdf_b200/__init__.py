"""B200-native dual-TSDF fusion hot path (drop-in for mast3r_slam/tsdf of
MASt3R-SLAM-Quality-dualTSDF).  CUDA kernels live in csrc/, reached through
the C ABI of include/b200tsdf.h via _ffi.py; there is no CPU fallback."""
__version__ = "0.1.0"
