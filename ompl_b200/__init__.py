"""ompl_b200 -- B200-native hot path of OMPL 1.6.0 (k-nearest / radius-nearest
queries, batched state validity, batched motion validity) behind the reference's
plug-in interfaces.  The compute lives in lib/libomplcu.so (hand-written CUDA
for sm_100a, C ABI in include/omplcu.h); this package is the host-side mirror.
"""
from ._lib import LIB_PATH, OmplCuError, lib  # noqa: F401
from .api import (  # noqa: F401
    AllValidStateValidityChecker,
    BatchedDiscreteMotionValidator,
    Circles2DValidityChecker,
    CompoundStateSpace,
    Exception_,
    HypercubeValidityChecker,
    KinematicChainSpace,
    KinematicChainValidityChecker,
    NearestNeighborsCUDA,
    RealVectorStateSpace,
    SE2StateSpace,
    SE3ObstacleValidityChecker,
    SE3StateSpace,
    SO2StateSpace,
    SO3StateSpace,
    StateSpace,
    StateValidityChecker,
)
