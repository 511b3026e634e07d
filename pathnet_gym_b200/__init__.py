"""pathnet_gym_b200 -- B200-native hot path of jan-hofmeier/pathnet-gym.

Host-side mirror of the reference's Python interface (pathnet, game_ac_network,
a3c_training_thread, rmsprop_applier, constants) over libpathnet_b200.so
(hand-written sm_100a CUDA behind a C ABI, include/pathnet_b200.h).
"""
from ._native import PathNetError, load as load_library  # noqa: F401

__all__ = ["PathNetError", "load_library"]
