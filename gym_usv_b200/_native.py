"""ctypes binding of libusvgpu.so (the C ABI declared in include/usvgpu.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is present the
calls raise.  ``build()`` compiles the library in-tree with nvcc for sm_100a.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.path.join(_PKG, "libusvgpu.so")
CSRC = os.path.join(_PKG, "csrc")
SOURCES = [os.path.join(CSRC, "usvgpu.cu")]
HEADERS = [os.path.join(CSRC, f) for f in ("usv_legacy.cuh", "usv_ca.cuh", "usv_render.cuh", "usv_info.cuh", "usv_model.cuh", "usv_philox.cuh", "usv_state.cuh", "usv_reset.cuh",
                                             "usv_sensor.cuh", "usv_kernels.cuh", "usv_math.cuh")] + \
          [os.path.join(_ROOT, "include", "usvgpu.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC"]

INFO_DIM = 24
# UsvInfoCol (include/usvgpu.h): key of the reference's info dict -> (first column, width)
INFO_KEYS = {"position": (0, 3), "velocity": (3, 3), "path_start": (6, 2), "path_end": (8, 2), "action0": (10, 1),
             "action1": (11, 1), "ye": (12, 1), "angle_to_target": (13, 1), "ye_reward": (14, 1),
             "angle_to_target_reward": (15, 1), "angle_action_reward": (16, 1), "delta_action_reward": (17, 1),
             "delta_action": (18, 1), "velocity_track_reward": (19, 1), "reference_velocity": (20, 1),
             "reward_velocity": (21, 1), "reference_velocity_error": (22, 1)}
OBS_DIM = 143
ACT_DIM = 2
SENSORS = 128
MAX_OBSTACLES = 32
DYN_ROWS = 23
EPI_ROWS = 7

ENV_ASMC_SIMPLE = 0
ENV_ASMC_V0 = 1
ENV_SIMPLE = 2
ENV_ASMC_YE_INT_V0, ENV_PID_V0 = 3, 4
ENV_ASMC_CA_V0 = 5
PREC_F64 = 0
PREC_MIXED = 1

FIELD_DYN, FIELD_EPI, FIELD_OBSTACLES, FIELD_N_OBS, FIELD_STEPS, FIELD_EPISODE, FIELD_EP_RETURN = range(7)
FIELD_LEGACY, FIELD_LEGACY_TARGET, FIELD_LEGACY_POS0, FIELD_LEGACY64 = 7, 8, 9, 10
FIELD_CA, FIELD_CA_N_OBS = 11, 12
FIELD_TARGET0 = 13
LEGACY_ROWS = 19
LEGACY64_ROWS = 21
LEGACY_OBS_DIM = 6
CA_ROWS = 31
CA_OBS_DIM = 23
CA_MAX_OBSTACLES = 9
LEGACY_NAMES = ("action_last", "u", "v", "r", "x", "y", "psi", "e_u_int", "ka_u", "ka_psi", "x_dot_last", "y_dot_last",
                "psi_dot_last", "u_dot_last", "v_dot_last", "r_dot_last", "e_u_last", "ka_dot_u_last", "ka_dot_psi_last")

# row names of FIELD_DYN / FIELD_EPI (include/usvgpu.h)
DYN_NAMES = ("x", "y", "psi", "u", "v", "r", "x_dot_last", "y_dot_last", "psi_dot_last", "u_dot_last", "v_dot_last",
             "r_dot_last", "e_u_last", "ka_dot_u_last", "ka_dot_psi_last", "psi_d_last", "o_dot_dot_last",
             "o_dot_last", "o_last", "e_u_int", "ka_u", "ka_psi", "progress")
EPI_NAMES = ("path_start_x", "path_start_y", "path_end_x", "path_end_y", "max_action_0", "max_action_2",
             "reference_velocity")


class UsvError(RuntimeError):
    pass


class UsvConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("env_kind", "ignore_obstacles", "do_perturb", "skip_kinematic",
                                         "max_episode_steps", "auto_reset", "precision", "obstacles_f32")]


class UsvStats(C.Structure):
    _fields_ = [("return_sum", C.c_double), ("length_sum", C.c_double), ("episodes", C.c_uint64),
                ("terminated", C.c_uint64), ("out_of_bounds", C.c_uint64), ("time_limit", C.c_uint64),
                ("env_steps", C.c_uint64), ("nonfinite", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


HASH_PATH = LIB_PATH + ".hash"   # sidecar: the source hash the binary was built from (also compiled into it)


def source_hash() -> str:
    """sha256 (first 16 hex digits) over the CUDA sources and headers of this tree: `usvgpu_source_hash()` of the
    loaded library must return the same string, which makes "the tested binary is this tree" checkable."""
    h = hashlib.sha256()
    for path in sorted(SOURCES + HEADERS):
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return nvcc if os.path.exists(nvcc) else None


def _stale() -> bool:
    """True when libusvgpu.so is missing or was built from other sources than the ones in this tree."""
    if not os.path.exists(LIB_PATH) or not os.path.exists(HASH_PATH):
        return True
    try:
        with open(HASH_PATH) as f:
            return f.read().strip() != source_hash()
    except OSError:
        return True


def build(force: bool = False, verbose: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> gym_usv_b200/libusvgpu.so (in-tree)."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    if nvcc is None:
        raise UsvError("nvcc not found and libusvgpu.so is missing or stale: cannot build the CUDA library")
    digest = source_hash()
    cmd = [nvcc] + NVCC_FLAGS + [f'-DUSVGPU_SOURCE_HASH="{digest}"'] + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB_PATH] + SOURCES
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise UsvError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    with open(HASH_PATH, "w") as f:
        f.write(digest + "\n")
    return LIB_PATH


_lib = None

_SIGNATURES = {
    "usvgpu_version": (C.c_int, []),
    "usvgpu_source_hash": (C.c_char_p, []),
    "usvgpu_last_error": (C.c_char_p, [C.c_void_p]),
    "usvgpu_create": (C.c_int, [C.POINTER(UsvConfig), C.c_int, C.c_int64, C.c_int64, C.c_uint64, C.POINTER(C.c_void_p)]),
    "usvgpu_destroy": (C.c_int, [C.c_void_p]),
    "usvgpu_obs_dim": (C.c_int, [C.c_void_p]),
    "usvgpu_act_dim": (C.c_int, [C.c_void_p]),
    "usvgpu_reset": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_step": (C.c_int, [C.c_void_p] * 8),
    "usvgpu_step_timed": (C.c_int, [C.c_void_p] * 9),
    "usvgpu_rollout": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p]),
    "usvgpu_use_staging": (C.c_int, [C.c_void_p] * 6),
    "usvgpu_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_int64]),
    "usvgpu_host_free": (C.c_int, [C.c_void_p]),
    "usvgpu_reset_host": (C.c_int, [C.c_void_p] * 4),
    "usvgpu_step_host": (C.c_int, [C.c_void_p] * 6 + [C.c_int32, C.c_void_p]),
    "usvgpu_stream_sync": (C.c_int, [C.c_void_p, C.c_void_p]),
    "usvgpu_set_host_chunks": (C.c_int, [C.c_void_p, C.c_int32]),
    "usvgpu_flags_sync": (C.c_int, [C.c_void_p]),
    "usvgpu_terminal_obs_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "usvgpu_field_bytes": (C.c_int64, [C.c_void_p, C.c_int32]),
    "usvgpu_get_state": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "usvgpu_set_state": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "usvgpu_episode_stats": (C.c_int, [C.c_void_p, C.POINTER(UsvStats), C.c_int32, C.c_void_p]),
    "usvgpu_stats_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "usvgpu_asmc_compute": (C.c_int, [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int64,
                                      C.c_int32, C.c_void_p, C.c_void_p]),
    "usvgpu_sensors": (C.c_int, [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_sensors_noalloc": (C.c_int, [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_frame_stack_push": (C.c_int, [C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 7),
    "usvgpu_render": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "usvgpu_step_info": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "usvgpu_reset_obs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_reset_ca_options": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_replay_add": (C.c_int, [C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int64, C.c_void_p, C.c_int64,
                                    C.c_void_p, C.c_int64, C.c_void_p, C.c_int64] + [C.c_void_p] * 11),
    "usvgpu_replay_sample": (C.c_int, [C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int64] + [C.c_void_p] * 14),
    "usvgpu_gae": (C.c_int, [C.c_int, C.c_int32, C.c_int64] + [C.c_void_p] * 5 + [C.c_double, C.c_double] + [C.c_void_p] * 3),
    "usvgpu_math_probe": (C.c_int, [C.c_int, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "usvgpu_fp64_peak": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.c_void_p]),
}

EXPORTS = tuple(_SIGNATURES)
CA_MAX_OBSTACLES = 9  # USVGPU_CA_MAX_OBSTACLES
NOALLOC_MAX_ENVS = 8192  # USVGPU_NOALLOC_MAX_ENVS


class UsvCaResetOptions(C.Structure):  # include/usvgpu.h
    _fields_ = [("has_obstacles", C.c_int32), ("n_obstacles", C.c_int32), ("obs_x", C.c_double * CA_MAX_OBSTACLES),
                ("obs_y", C.c_double * CA_MAX_OBSTACLES), ("obs_r", C.c_double * CA_MAX_OBSTACLES),
                ("has_target", C.c_int32), ("has_start", C.c_int32), ("target", C.c_double * 2), ("start", C.c_double * 2)]


def ca_reset_options(options: dict) -> UsvCaResetOptions:
    """The option dict of UsvAsmcCaEnv.reset (gym_usv/envs/usv_asmc_ca_env.py:358-372: 'obs_x' / 'obs_y' / 'obs_r',
    'target_point', 'start_position'; 'renderplots' belongs to the reference's plotting and is ignored) as the C struct."""
    import numpy as np
    unknown = set(options) - {"obs_x", "obs_y", "obs_r", "target_point", "start_position", "renderplots"}
    if unknown:
        raise UsvError(f"usv-asmc-ca-v0 reset options: unknown key(s) {sorted(unknown)}")
    opt = UsvCaResetOptions()
    if "obs_x" in options:
        try:
            ox, oy, orr = (np.asarray(options[k], dtype=np.float64).reshape(-1) for k in ("obs_x", "obs_y", "obs_r"))
        except KeyError as e:
            raise UsvError(f"reset options: 'obs_x' needs 'obs_y' and 'obs_r' beside it (missing {e})") from None
        if not (len(ox) == len(oy) == len(orr)):
            raise UsvError("reset options: obs_x / obs_y / obs_r differ in length")
        if len(ox) > CA_MAX_OBSTACLES:
            raise UsvError(f"reset options: {len(ox)} obstacles; this build holds {CA_MAX_OBSTACLES} per env")
        opt.has_obstacles, opt.n_obstacles = 1, len(ox)
        for j in range(len(ox)):
            opt.obs_x[j], opt.obs_y[j], opt.obs_r[j] = float(ox[j]), float(oy[j]), float(orr[j])
    for key, flag, dst in (("target_point", "has_target", opt.target), ("start_position", "has_start", opt.start)):
        if key in options:
            v = np.asarray(options[key], dtype=np.float64).reshape(-1)
            if len(v) < 2:
                raise UsvError(f"reset options: '{key}' needs at least x and y")
            setattr(opt, flag, 1)
            dst[0], dst[1] = float(v[0]), float(v[1])
    return opt


def lib():
    """Load libusvgpu.so.  A binary that is missing or was built from other sources than this tree's is rebuilt when
    nvcc is here; without nvcc a stale binary is an error, not a silent mismatch (USVGPU_LIB, an alternative build of
    the same ABI for A/B experiments, is loaded as it is)."""
    global _lib
    if _lib is None:
        path = os.environ.get("USVGPU_LIB", LIB_PATH)
        if path == LIB_PATH and _stale():
            if _nvcc() is None:
                raise UsvError("libusvgpu.so is missing or was not built from this tree's sources, and nvcc is not "
                               "here to rebuild it (python -c 'import __graft_entry__ as g; g.build()')")
            build()
        L = C.CDLL(path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        if path == LIB_PATH and L.usvgpu_source_hash().decode() != source_hash():
            raise UsvError("libusvgpu.so reports another source hash than this tree's: rebuild it")
        _lib = L
    return _lib


def device_index(device) -> int:
    """CUDA ordinal of a torch device; a bare "cuda" (index None) means the CURRENT device, not device 0."""
    if device.index is not None:
        return int(device.index)
    import torch
    return int(torch.cuda.current_device())


def check(rc: int, handle=None) -> None:
    if rc != 0:
        msg = lib().usvgpu_last_error(handle)
        raise UsvError(f"libusvgpu error {rc}: {msg.decode() if msg else '?'}")
