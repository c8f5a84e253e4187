"""Batched stand-alone entry points (unit-parity surface of the inner reference API)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as nat


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def asmc_compute(state: torch.Tensor, actions: torch.Tensor, n_calls: int = 1, do_perturb: bool = False,
                 first_substep: int = 0, precision: str = "f64", want_info: bool = False):
    """UsvAsmc.compute (gym_usv/control/usv_asmc.py:53-244) for n independent states, in place.

    state   float64 CUDA [23, n]  rows as USV_FIELD_DYN (x y psi u v r | last[9] | psi_d_last o_dot_dot_last
            o_dot_last o_last | e_u_int ka_u ka_psi | unused)
    actions float64 CUDA [n, 2]
    Returns info float64 [4, n] (psi_d, u_d, tport, tstbd of the last sub-step) when want_info.
    """
    if not (state.is_cuda and actions.is_cuda):
        raise nat.UsvError("asmc_compute needs CUDA tensors: there is no CPU fallback")
    assert state.dtype == torch.float64 and state.is_contiguous() and state.shape[0] == nat.DYN_ROWS
    n = state.shape[1]
    assert actions.dtype == torch.float64 and actions.is_contiguous() and tuple(actions.shape) == (n, 2)
    info = torch.zeros((4, n), dtype=torch.float64, device=state.device) if want_info else None
    rc = nat.lib().usvgpu_asmc_compute(nat.device_index(state.device), n, C.c_void_p(state.data_ptr()),
                                       C.c_void_p(actions.data_ptr()), int(n_calls), int(do_perturb),
                                       int(first_substep), {"f64": 0, "mixed": 1}[precision],
                                       C.c_void_p(info.data_ptr()) if want_info else None, _stream(state.device))
    nat.check(rc)
    return info


def sensor_distances(pose: torch.Tensor, obstacles: torch.Tensor, n_obs: torch.Tensor) -> torch.Tensor:
    """UsvAsmcCaEnv.compute_sensor_measurments (usv_asmc_ca_env.py:411-437) with the 128-ray fan of
    simple_env.py:11-14.  pose float64 [3, n]; obstacles float64 [32, 3, n]; n_obs int32 [n].
    Returns distances float64 [128, n]."""
    if not pose.is_cuda:
        raise nat.UsvError("sensor_distances needs CUDA tensors: there is no CPU fallback")
    n = pose.shape[1]
    assert pose.dtype == torch.float64 and pose.is_contiguous() and pose.shape[0] == 3
    assert obstacles.dtype == torch.float64 and obstacles.is_contiguous() and tuple(obstacles.shape) == (32, 3, n)
    assert n_obs.dtype == torch.int32 and n_obs.is_contiguous() and n_obs.shape[0] == n
    out = torch.empty((nat.SENSORS, n), dtype=torch.float64, device=pose.device)
    rc = nat.lib().usvgpu_sensors(nat.device_index(pose.device), n, C.c_void_p(pose.data_ptr()),
                                  C.c_void_p(obstacles.data_ptr()), C.c_void_p(n_obs.data_ptr()),
                                  C.c_void_p(out.data_ptr()), _stream(pose.device))
    nat.check(rc)
    return out


def sensor_distances_noalloc(pose: torch.Tensor, obstacles: torch.Tensor, n_obs: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """sensor_distances through usvgpu_sensors_noalloc (no counterpart in the reference; include/usvgpu.h): the library
    allocates nothing -- its scratch is a fixed static device array -- and the caller brings ``out`` (float64 [128, n]).
    At most ``NOALLOC_MAX_ENVS`` envs per call; calls that overlap on the GPU race on the shared scratch."""
    if not pose.is_cuda:
        raise nat.UsvError("sensor_distances_noalloc needs CUDA tensors: there is no CPU fallback")
    n = pose.shape[1]
    assert pose.dtype == torch.float64 and pose.is_contiguous() and pose.shape[0] == 3
    assert obstacles.dtype == torch.float64 and obstacles.is_contiguous() and tuple(obstacles.shape) == (32, 3, n)
    assert n_obs.dtype == torch.int32 and n_obs.is_contiguous() and n_obs.shape[0] == n
    assert out.dtype == torch.float64 and out.is_contiguous() and tuple(out.shape) == (nat.SENSORS, n) and out.device == pose.device
    nat.check(nat.lib().usvgpu_sensors_noalloc(nat.device_index(pose.device), n, C.c_void_p(pose.data_ptr()),
                                               C.c_void_p(obstacles.data_ptr()), C.c_void_p(n_obs.data_ptr()),
                                               C.c_void_p(out.data_ptr()), _stream(pose.device)))
    return out


def math_probe(which: str, x: torch.Tensor) -> torch.Tensor:
    """Evaluate one float64 device routine of csrc/usv_math.cuh (test surface).  which: 'sqrt' / 'sqrt_m' x[n] | 'div' x[2,n] |
    'asin' x[n] | 'rot' x[3,n] = (delta, cos, sin) -> [2,n]."""
    code = {"sqrt": 0, "div": 1, "asin": 2, "rot": 3, "sqrt_m": 4}[which]
    if not x.is_cuda:
        raise nat.UsvError("math_probe needs CUDA tensors: there is no CPU fallback")
    assert x.dtype == torch.float64 and x.is_contiguous()
    n = x.shape[-1]
    out = torch.empty((2, n) if which == "rot" else (n,), dtype=torch.float64, device=x.device)
    nat.check(nat.lib().usvgpu_math_probe(nat.device_index(x.device), code, n, C.c_void_p(x.data_ptr()),
                                          C.c_void_p(out.data_ptr()), _stream(x.device)))
    return out


def fp64_peak(device: int = 0, blocks: int = 148 * 8, iters: int = 20000) -> float:
    """Measured thread-level float64 FMA/s of the GPU (usvgpu_fp64_peak)."""
    v = C.c_double()
    nat.check(nat.lib().usvgpu_fp64_peak(int(device), int(blocks), int(iters), C.byref(v),
                                         _stream(torch.device("cuda", device))))
    return float(v.value)
