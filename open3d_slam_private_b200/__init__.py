"""B200-native point-cloud registration core (hot path of leggedrobotics/open3d_slam_private).

Only the ICP hot path lives here: csrc/ (CUDA kernels for sm_100a + the C ABI of include/b200reg.h), capi.py (ctypes
binding), pointmatcher.py (host-side mirror of the reference's module interface) and synth.py (synthetic LiDAR
workloads for the benchmark). The CUDA library is mandatory: nothing in this package computes on the CPU.
"""
from . import capi  # noqa: F401

__all__ = ["capi"]
