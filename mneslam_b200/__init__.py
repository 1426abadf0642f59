"""mneslam_b200: B200-native (sm_100a) implementation of MNE-SLAM's JointEncoding ray-batch hot path."""
from .scene_rep import JointEncoding  # noqa: F401

__all__ = ['JointEncoding']
