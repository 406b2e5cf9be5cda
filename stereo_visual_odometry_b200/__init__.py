"""B200-native stereo visual-odometry front end (sm_100a CUDA kernels behind the
reference's include/vo.h interface).  See DESIGN.md / INTEGRATION.md."""
from ._lib import FrameResult, Params, SvoError, build, default_params, lib  # noqa: F401
from .vo import (FeatureSet, VisualOdometry, cameraToWorld, circularMatching,  # noqa: F401
                 deleteFeaturesWithFailureStatus, deletePointsWithFailureStatus, featureDetectionFast,
                 findClosePoints, getInverseTransform, triangulatePoints)
