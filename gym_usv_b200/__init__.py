"""gym_usv_b200 -- B200-native batched env step for romi2002/gym-usv's ASMC path-following env.

Public surface (mirrors what the reference's callers use, train_test/sb3_train_vec.py:67-81):
  UsvGpuVecEnv   SB3-style VecEnv for 'usv-asmc-simple' (host numpy face + device tensor face)
  make_vec_env / UsvGpuSb3VecEnv / register_envs / UsvGpuEnv   the seats of SB3's make_vec_env, of BaseAlgorithm's VecEnv
                 check, of the gymnasium registry and of gym.make callers (sb3.py)
  asmc_compute   batched UsvAsmc.compute (gym_usv/control/usv_asmc.py:53-244)
  sensor_distances  batched UsvAsmcCaEnv.compute_sensor_measurments (usv_asmc_ca_env.py:411-437)
The arithmetic lives in libusvgpu.so (csrc/, hand-written sm_100a CUDA behind include/usvgpu.h).
"""
from ._native import UsvError, build  # noqa: F401
from .vec_env import UsvGpuVecEnv  # noqa: F401
from .frame_stack import UsvGpuVecFrameStack  # noqa: F401
from .replay import DeviceReplayBuffer, collect_transitions  # noqa: F401
from .video import UsvGpuVecVideoRecorder, write_apng  # noqa: F401
from .functional import asmc_compute, fp64_peak, math_probe, sensor_distances  # noqa: F401
from .sb3 import UsvGpuEnv, UsvGpuSb3VecEnv, make_vec_env, register_envs, sb3_vec_env_class  # noqa: F401

__all__ = ["UsvGpuEnv", "UsvGpuSb3VecEnv", "make_vec_env", "register_envs", "sb3_vec_env_class", "UsvGpuVecEnv", "UsvGpuVecFrameStack", "DeviceReplayBuffer", "collect_transitions", "UsvGpuVecVideoRecorder", "write_apng", "UsvError", "build", "asmc_compute", "sensor_distances", "math_probe", "fp64_peak"]
