"""ctypes binding of the CPU oracle (oracle/libusv_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py -- never by gym_usv_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libusv_oracle.so")

MAX_OBS = 32
SENSORS = 128
OBS_DIM = 143

ASMC_DTYPE = np.dtype(
    [
        ("eta", "<f8", (3,)),
        ("nu", "<f8", (3,)),
        ("last", "<f8", (9,)),
        ("so", "<f8", (4,)),
        ("aux", "<f8", (3,)),
        ("perturb_step", "<i8"),
    ],
    align=True,
)

ENV_DTYPE = np.dtype(
    [
        ("asmc", ASMC_DTYPE),
        ("path_start", "<f8", (2,)),
        ("path_end", "<f8", (2,)),
        ("target", "<f8", (2,)),
        ("max_action", "<f8", (3,)),
        ("max_accel", "<f8", (3,)),
        ("v_ref", "<f8"),
        ("progress", "<f8"),
        ("last_action", "<f8", (3,)),
        ("obs_x", "<f8", (MAX_OBS,)),
        ("obs_y", "<f8", (MAX_OBS,)),
        ("obs_r", "<f8", (MAX_OBS,)),
        ("sensor", "<f8", (SENSORS,)),
        ("ep_return", "<f8"),
        ("n_obs", "<i4"),
        ("steps", "<i4"),
        ("episode", "<i4"),
        ("pad_", "<i4"),
    ],
    align=True,
)


LEGACY_DTYPE = np.dtype(
    [
        ("state", "<f8", (6,)),
        ("velocity", "<f8", (3,)),
        ("position", "<f8", (3,)),
        ("aux", "<f8", (3,)),
        ("last", "<f8", (9,)),
        ("target", "<f8", (6,)),
        ("ep_return", "<f8"),
        ("ye_int", "<f8"),
        ("ye_last", "<f8"),
        ("f32_state", "<i4"),
        ("steps", "<i4"),
        ("episode", "<i4"),
        ("variant", "<i4"),
    ],
    align=True,
)
CA_OBS_DIM, CA_MAX_OBS = 23, 9
CA_DTYPE = np.dtype(
    [
        ("asmc", ASMC_DTYPE),
        ("target", "<f8", (2,)),
        ("prev_action", "<f8", (2,)),
        ("obs_x", "<f8", (CA_MAX_OBS,)),
        ("obs_y", "<f8", (CA_MAX_OBS,)),
        ("obs_r", "<f8", (CA_MAX_OBS,)),
        ("ep_return", "<f8"),
        ("n_obs", "<i4"),
        ("steps", "<i4"),
        ("episode", "<i4"),
        ("pad_", "<i4"),
    ],
    align=True,
)
# gym_usv/__init__.py:3-16: the three registered ids that share UsvAsmcEnv's step()
LEGACY_VARIANTS = {"usv-asmc-v0": 0, "usv-asmc-ye-int-v0": 1, "usv-pid-v0": 2}


class Config(C.Structure):
    _fields_ = [
        ("ignore_obstacles", C.c_int32),
        ("do_perturb", C.c_int32),
        ("skip_kinematic", C.c_int32),
        ("max_episode_steps", C.c_int32),
        ("auto_reset", C.c_int32),
        ("obstacles_f32", C.c_int32),
        ("n_threads", C.c_int32),
        ("kinematic", C.c_int32),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (make -C oracle).  Returns the .so path."""
    srcs = [os.path.join(_HERE, f) for f in ("usv_oracle.c", "usv_oracle_legacy.c", "usv_oracle_legacy_step.inc", "usv_oracle_render.c", "usv_oracle_ca.c",
                                             "usv_oracle.h")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs if os.path.exists(s)
    )
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
        L.usv_oracle_sizeof_env.restype = C.c_int
        L.usv_oracle_sizeof_asmc.restype = C.c_int
        L.usv_oracle_asmc_init.argtypes = [vp]
        L.usv_oracle_asmc_compute.argtypes = [vp, vp, i32, vp]
        L.usv_oracle_asmc_compute_batch.argtypes = [vp, i64, vp, i32, i32, i32]
        L.usv_oracle_sensors.argtypes = [vp, i32, dbl, vp, dbl, vp, vp, i32, vp, vp]
        L.usv_oracle_simple_init.argtypes = [vp]
        L.usv_oracle_simple_reset.argtypes = [vp, vp, u64, u64, vp]
        L.usv_oracle_simple_reset_obs.argtypes = [vp, vp]
        L.usv_oracle_simple_reset_obs.restype = None
        L.usv_oracle_simple_step.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.usv_oracle_vec_reset.argtypes = [vp, i64, vp, u64, i64, vp]
        L.usv_oracle_vec_step.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, vp, u64, i64]
        L.usv_oracle_philox.argtypes = [u64, u64, C.c_uint32, C.c_uint32, vp]
        for name in ("usv_oracle_asmc_init", "usv_oracle_asmc_compute", "usv_oracle_asmc_compute_batch",
                     "usv_oracle_sensors", "usv_oracle_simple_init", "usv_oracle_simple_reset",
                     "usv_oracle_simple_step", "usv_oracle_vec_reset", "usv_oracle_vec_step",
                     "usv_oracle_philox"):
            getattr(L, name).restype = None
        L.usv_oracle_sizeof_legacy.restype = C.c_int
        L.usv_oracle_legacy_init.argtypes = [vp]
        L.usv_oracle_legacy_reset_from_draws.argtypes = [vp, vp, vp]
        L.usv_oracle_legacy_reset.argtypes = [vp, u64, u64, vp]
        L.usv_oracle_legacy_step.argtypes = [vp, C.c_float, vp, vp, vp]
        L.usv_oracle_legacy_vec_reset.argtypes = [vp, i64, u64, i64, vp, i32]
        L.usv_oracle_legacy_vec_step.argtypes = [vp, i64, vp, vp, vp, vp, vp, i32, u64, i64, i32]
        for name in ("usv_oracle_legacy_init", "usv_oracle_legacy_reset_from_draws", "usv_oracle_legacy_reset",
                     "usv_oracle_legacy_step", "usv_oracle_legacy_vec_reset", "usv_oracle_legacy_vec_step"):
            getattr(L, name).restype = None
        L.usv_oracle_sizeof_ca.restype = C.c_int
        L.usv_oracle_ca_vec_reset.argtypes = [vp, i64, u64, i64, vp, i32]
        L.usv_oracle_ca_vec_step.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, i32, i32, u64, i64, i32]
        L.usv_oracle_ca_vec_reset_opt.argtypes = [vp, i64, u64, i64, vp, vp, i32]
        L.usv_oracle_ca_vec_reset.restype = L.usv_oracle_ca_vec_step.restype = L.usv_oracle_ca_vec_reset_opt.restype = None
        assert L.usv_oracle_sizeof_ca() == CA_DTYPE.itemsize, (L.usv_oracle_sizeof_ca(), CA_DTYPE.itemsize)
        L.usv_oracle_render_frame.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, dbl, i32, vp]
        L.usv_oracle_render_frame.restype = None
        assert L.usv_oracle_sizeof_legacy() == LEGACY_DTYPE.itemsize
        assert L.usv_oracle_sizeof_env() == ENV_DTYPE.itemsize, (L.usv_oracle_sizeof_env(), ENV_DTYPE.itemsize)
        assert L.usv_oracle_sizeof_asmc() == ASMC_DTYPE.itemsize
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_config(ignore_obstacles=False, do_perturb=False, skip_kinematic=False, max_episode_steps=1000,
                auto_reset=True, obstacles_f32=False, n_threads=0, kinematic=False) -> Config:
    return Config(int(ignore_obstacles), int(do_perturb), int(skip_kinematic), int(max_episode_steps),
                  int(auto_reset), int(obstacles_f32), int(n_threads), int(kinematic))


# ---------------------------------------------------------------------------------------
# UsvAsmc.compute (gym_usv/control/usv_asmc.py:53-244)
# ---------------------------------------------------------------------------------------
class OracleAsmc:
    """Batch of N controller+vehicle states."""

    def __init__(self, n=1):
        self.s = np.zeros(n, dtype=ASMC_DTYPE)

    def compute(self, actions, do_perturb=False, n_calls=1, n_threads=1):
        a = np.ascontiguousarray(np.broadcast_to(np.asarray(actions, dtype=np.float64), (len(self.s), 2)))
        lib().usv_oracle_asmc_compute_batch(_p(self.s), len(self.s), _p(a), int(do_perturb), int(n_calls),
                                            int(n_threads))
        return self.s["eta"].copy(), self.s["nu"].copy()

    def compute_one(self, i, action, do_perturb=False):
        a = np.asarray(action, dtype=np.float64)
        info = np.zeros(4)
        lib().usv_oracle_asmc_compute(self.s[i:i + 1].ctypes.data_as(C.c_void_p), _p(a), int(do_perturb), _p(info))
        return info


def sensors(position, radius, posx, posy, sensor_count=SENSORS, max_range=100.0, pitch=None):
    """UsvAsmcCaEnv.compute_sensor_measurments (usv_asmc_ca_env.py:411-437) -> distances[S]."""
    position = np.ascontiguousarray(position, dtype=np.float64)
    radius = np.ascontiguousarray(radius, dtype=np.float64)
    posx = np.ascontiguousarray(posx, dtype=np.float64).reshape(-1)
    posy = np.ascontiguousarray(posy, dtype=np.float64).reshape(-1)
    if pitch is None:
        pitch = ((2 / 3) * (2 * np.pi)) / sensor_count
    dist = np.hypot(posx - position[0], posy - position[1]) - radius
    out = np.zeros(sensor_count)
    lib().usv_oracle_sensors(_p(position), sensor_count, float(max_range), _p(radius), float(pitch), _p(posx),
                             _p(posy), len(radius), _p(dist), _p(out))
    return out


# ---------------------------------------------------------------------------------------
# usv-asmc-simple vector env (simple_env_asmc.py + simple_env.py + TimeLimit)
# ---------------------------------------------------------------------------------------
class OracleVecEnv:
    def __init__(self, n, seed=0, global_offset=0, **cfg):
        self.n, self.seed, self.global_offset = int(n), int(seed), int(global_offset)
        self.cfg = make_config(**cfg)
        self.envs = np.zeros(self.n, dtype=ENV_DTYPE)
        for i in range(self.n):
            lib().usv_oracle_simple_init(self.envs[i:i + 1].ctypes.data_as(C.c_void_p))
        self.obs = np.zeros((self.n, OBS_DIM), dtype=np.float32)
        self.terminal_obs = np.zeros((self.n, OBS_DIM), dtype=np.float32)
        self.reward = np.zeros(self.n)
        self.terminated = np.zeros(self.n, dtype=np.uint8)
        self.truncated = np.zeros(self.n, dtype=np.uint8)

    def reset(self):
        lib().usv_oracle_vec_reset(_p(self.envs), self.n, C.byref(self.cfg), self.seed, self.global_offset,
                                   _p(self.obs))
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32).reshape(self.n, 2)
        lib().usv_oracle_vec_step(_p(self.envs), self.n, C.byref(self.cfg), _p(a), _p(self.obs), _p(self.reward),
                                  _p(self.terminated), _p(self.truncated), _p(self.terminal_obs), self.seed,
                                  self.global_offset)
        return self.obs, self.reward, self.terminated, self.truncated


# ---------------------------------------------------------------------------------------
# legacy path-following vector envs: usv-asmc-v0 (gym_usv/envs/usv_asmc_env.py) and its two float64 siblings
# usv-asmc-ye-int-v0 (usv_asmc_ye_int_env.py) and usv-pid-v0 (usv_pid_env.py)
# ---------------------------------------------------------------------------------------
class OracleLegacyVecEnv:
    """N UsvAsmcEnv / UsvAsmcYeIntEnv / UsvPidEnv instances in lock-step, SB3 auto-reset semantics."""

    def __init__(self, n, seed=0, global_offset=0, auto_reset=True, n_threads=1, env_id="usv-asmc-v0"):
        self.n, self.seed, self.global_offset = int(n), int(seed), int(global_offset)
        self.auto_reset, self.n_threads = bool(auto_reset), int(n_threads)
        self.envs = np.zeros(self.n, dtype=LEGACY_DTYPE)
        lib()
        self.envs["episode"] = -1
        self.envs["variant"] = LEGACY_VARIANTS[env_id]
        self.obs = np.zeros((self.n, 6), dtype=np.float32)
        self.terminal_obs = np.zeros((self.n, 6), dtype=np.float32)
        self.reward = np.zeros(self.n)
        self.done = np.zeros(self.n, dtype=np.uint8)

    def reset(self):
        lib().usv_oracle_legacy_vec_reset(_p(self.envs), self.n, self.seed, self.global_offset, _p(self.obs),
                                          self.n_threads)
        return self.obs

    def reset_from_draws(self, draws):
        """usv_asmc_env.py:258-300 from the reference's own seven uniform draws per env."""
        d = np.ascontiguousarray(draws, dtype=np.float64).reshape(self.n, 7)
        for i in range(self.n):
            lib().usv_oracle_legacy_reset_from_draws(self.envs[i:i + 1].ctypes.data_as(C.c_void_p), _p(d[i]),
                                                     _p(self.obs[i]))
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32).reshape(self.n)
        lib().usv_oracle_legacy_vec_step(_p(self.envs), self.n, _p(a), _p(self.obs), _p(self.reward), _p(self.done),
                                         _p(self.terminal_obs), int(self.auto_reset), self.seed, self.global_offset,
                                         self.n_threads)
        return self.obs, self.reward, self.done


class CaOptions(C.Structure):  # UsvOracleCaOptions, usv_oracle.h
    _fields_ = [("has_obstacles", C.c_int32), ("n_obstacles", C.c_int32), ("obs_x", C.c_double * 9),
                ("obs_y", C.c_double * 9), ("obs_r", C.c_double * 9), ("has_target", C.c_int32), ("has_start", C.c_int32),
                ("target", C.c_double * 2), ("start", C.c_double * 2)]


def ca_options(options) -> CaOptions:
    opt = CaOptions()
    if "obs_x" in options:  # usv_asmc_ca_env.py:361-365
        ox, oy, orr = (np.asarray(options[k], dtype=np.float64).reshape(-1) for k in ("obs_x", "obs_y", "obs_r"))
        assert len(ox) == len(oy) == len(orr) <= 9
        opt.has_obstacles, opt.n_obstacles = 1, len(ox)
        opt.obs_x[:len(ox)], opt.obs_y[:len(ox)], opt.obs_r[:len(ox)] = ox.tolist(), oy.tolist(), orr.tolist()
    if "target_point" in options:
        opt.has_target, opt.target[:] = 1, [float(v) for v in options["target_point"][:2]]
    if "start_position" in options:
        opt.has_start, opt.start[:] = 1, [float(v) for v in options["start_position"][:2]]
    return opt


class OracleCaVecEnv:
    """N UsvAsmcCaEnv instances ('usv-asmc-ca-v0', in-repo UsvAsmc substituted for usv_libs_py) in lock-step,
    SB3 auto-reset semantics; TimeLimit 5000 (gym_usv/__init__.py:18-22)."""

    def __init__(self, n, seed=0, global_offset=0, auto_reset=True, max_episode_steps=5000, n_threads=1):
        self.n, self.seed, self.global_offset = int(n), int(seed), int(global_offset)
        self.auto_reset, self.max_episode_steps, self.n_threads = bool(auto_reset), int(max_episode_steps), int(n_threads)
        self.envs = np.zeros(self.n, dtype=CA_DTYPE)
        lib()
        self.envs["episode"] = -1
        self.obs = np.zeros((self.n, CA_OBS_DIM), dtype=np.float32)
        self.terminal_obs = np.zeros((self.n, CA_OBS_DIM), dtype=np.float32)
        self.reward = np.zeros(self.n)
        self.done = np.zeros(self.n, dtype=np.uint8)
        self.trunc = np.zeros(self.n, dtype=np.uint8)

    def reset(self, options=None):
        """options: the dict of UsvAsmcCaEnv.reset (usv_asmc_ca_env.py:358-372), applied to every env."""
        if options:
            opt = ca_options(options)
            lib().usv_oracle_ca_vec_reset_opt(_p(self.envs), self.n, self.seed, self.global_offset, C.addressof(opt),
                                              _p(self.obs), self.n_threads)
        else:
            lib().usv_oracle_ca_vec_reset(_p(self.envs), self.n, self.seed, self.global_offset, _p(self.obs), self.n_threads)
        return self.obs

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float32).reshape(self.n, 2)
        lib().usv_oracle_ca_vec_step(_p(self.envs), self.n, _p(a), _p(self.obs), _p(self.reward), _p(self.done),
                                     _p(self.trunc), _p(self.terminal_obs), int(self.auto_reset),
                                     self.max_episode_steps, self.seed, self.global_offset, self.n_threads)
        return self.obs, self.reward, self.done, self.trunc


def render_frame(position, target, sensor_distance, obs_x, obs_y, obs_r, path_start, path_end, env_hi=20.0, size=512):
    """SimpleEnvVisualizer.render_frame (simple_env_visualizer.py:17-114) -> uint8 [size, size, 3].  Parity unpinned
    against pygame's pixels (see usv_oracle_render.c)."""
    f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64).reshape(-1)  # noqa: E731
    position, target, d = f64(position), f64(target), f64(sensor_distance)
    ox, oy, orad, ps, pe = f64(obs_x), f64(obs_y), f64(obs_r), f64(path_start), f64(path_end)
    frame = np.zeros((size, size, 3), dtype=np.uint8)
    lib().usv_oracle_render_frame(_p(position), _p(target), _p(d), len(d), len(ox), _p(ox), _p(oy), _p(orad), _p(ps),
                                  _p(pe), float(env_hi), int(size), _p(frame))
    return frame


def philox(seed, gid, episode, block):
    out = np.zeros(4, dtype=np.uint32)
    lib().usv_oracle_philox(int(seed), int(gid), int(episode), int(block), _p(out))
    return out
