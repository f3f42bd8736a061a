"""B200-native batched scan core for ut_multirobot_sim's laser-scan hot path.

The product is libumrs.so (CUDA, sm_100a) behind the C-ABI in include/umrs.h;
this package is the thin Python host mirror used by tests and bench.py.
"""
from . import _lib  # noqa: F401
from .api import (Config, VectorMap, World, default_human, default_laser, default_robot,  # noqa: F401
                  laser_num_rays, map_cleanup)
from ._lib import ACKERMANN_DRIVE, DIFF_DRIVE, OMNIDIRECTIONAL_DRIVE  # noqa: F401
