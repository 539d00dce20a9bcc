"""adaptpoint_b200 -- B200-native (sm_100a) implementation of the PointNeXt set-abstraction hot path
of AdaptPoint/OpenPoints, behind the reference's operator API (openpoints/models/layers)."""
from .ops import (  # noqa: F401
    furthest_point_sample, gather_operation, ball_query, grouping_operation, three_nn, three_interpolate,
    three_interpolation, fps, knn_query, FurthestPointSampling, GatherOperation, BallQuery, GroupingOperation,
    ThreeNN, ThreeInterpolate,
)

__version__ = "0.1.0"
