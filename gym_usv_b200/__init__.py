"""gym_usv_b200 -- B200-native batched env step for romi2002/gym-usv's ASMC path-following env.

Public surface (mirrors what the reference's callers use, train_test/sb3_train_vec.py:67-81):
  UsvGpuVecEnv   SB3-style VecEnv for 'usv-asmc-simple' (host numpy face + device tensor face)
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

__all__ = ["UsvGpuVecEnv", "UsvGpuVecFrameStack", "DeviceReplayBuffer", "collect_transitions", "UsvGpuVecVideoRecorder", "write_apng", "UsvError", "build", "asmc_compute", "sensor_distances", "math_probe", "fp64_peak"]
