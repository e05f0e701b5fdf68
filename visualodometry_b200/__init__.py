"""visualodometry_b200 -- B200-native bundle adjustment for the cteufel13/VisualOdometry loop.

Only the BA hot path lives here: csrc/ (CUDA kernels + C ABI, libb2ba.so),
ba.py (host-side BundleAdjuster / BAConfig), adapter.py (VO keyframe-window adapter) and
synth.py (synthetic problems of the BASELINE.json shapes)."""
from .ba import BAConfig, BAResult, BASummary, BundleAdjuster, T_from_poses, poses_from_T  # noqa: F401

__all__ = ["BAConfig", "BAResult", "BASummary", "BundleAdjuster", "poses_from_T", "T_from_poses"]
