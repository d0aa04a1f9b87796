"""go2_b200 - B200-native batched Go2 locomotion environment (host-side mirror of the reference env API).

The hot path lives in csrc/ (hand-written CUDA for sm_100a) behind the C ABI in include/go2_b200.h.
"""
from .abi import CUDA_LIB_PATH, Go2Lib  # noqa: F401
from .envs import Go2EnvBaseFlavour, Go2EnvWalk  # noqa: F401
