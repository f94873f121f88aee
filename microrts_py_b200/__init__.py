"""microrts_py_b200: a B200-native (sm_100a CUDA) drop-in for the hot path of
adFrej/MicroRTS-Py -- the vectorised environment step behind ``MicroRTSGridModeVecEnv``.

Layout (mirrors ``gym_microrts``):
  microrts_py_b200.envs.MicroRTSGridModeVecEnv   the environment class (gym_microrts/envs/vec_env.py)
  microrts_py_b200.microrts_ai                   scripted opponents usable as ``ai2s`` entries
  microrts_py_b200.maps                          map XML loader + bundled map pack
  microrts_py_b200._capi                         ctypes binding of csrc/libmrts_b200.so (include/mrts_b200.h)

Importing the package does not import torch or load the CUDA library; constructing an
environment does, and raises if either the library or a CUDA device is missing.
"""

__version__ = "0.1.0"


def __getattr__(name):
    if name == "MicroRTSGridModeVecEnv":
        from .envs.vec_env import MicroRTSGridModeVecEnv

        return MicroRTSGridModeVecEnv
    raise AttributeError(name)
