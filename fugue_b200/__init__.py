"""fugue_b200 -- B200-native inference backend for fugue's HMC / MH / SMC hot path.

Host-side mirror of the reference's public entry points over the C ABI in
include/fugue_cuda.h (libfugue_cuda.so, hand-written sm_100a kernels). There is NO CPU
fallback: importing `fugue_b200.api` without the built extension raises.
"""
from .model import *  # noqa: F401,F403
from ._ir import FugueError  # noqa: F401
