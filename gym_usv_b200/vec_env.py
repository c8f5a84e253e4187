"""UsvGpuVecEnv -- drop-in SB3-style VecEnv for 'usv-asmc-simple' running on one B200.

Mirrors what ``make_vec_env('usv-asmc-simple', n_envs)`` gives the scripts in
train_test/sb3_train_vec.py:67-81: ``reset() -> obs``, ``step(actions) -> (obs, rewards, dones,
infos)`` with SB3 auto-reset semantics (``infos[i]['terminal_observation']``,
``['TimeLimit.truncated']``, ``['episode']``), same observation / action spaces
(gym_usv/envs/simple_env.py:27,30) and the 1000-step TimeLimit of gym_usv/__init__.py:30-34.

Two faces of the same env:
  * ``step(np_actions)``  -- host numpy in / host numpy out (what stock SB3 calls);
  * ``step_tensor(cuda_actions)`` -- device-resident torch tensors in and out, zero host copies.

All arithmetic happens in libusvgpu.so (hand-written sm_100a kernels).  torch is used only for
device memory and streams.  There is no CPU path: constructing the env without CUDA raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import _native as nat
from . import spaces

ENV_ID = "usv-asmc-simple"
# gym_usv/__init__.py:30-34 registers a 1000-step TimeLimit for 'usv-asmc-simple' and none for 'usv-asmc-v0' (:3-6)
# 'usv-pid-v0' (:8-11) and 'usv-asmc-ye-int-v0' (:13-16) are registered without a limit as well
MAX_EPISODE_STEPS = {"usv-asmc-simple": 1000, "usv-asmc-v0": 0, "usv-simple": 500,  # usv-simple: __init__.py:24-28
                     "usv-asmc-ye-int-v0": 0, "usv-pid-v0": 0, "usv-asmc-ca-v0": 5000}  # ca: __init__.py:18-22
ENV_KINDS = {"usv-asmc-simple": nat.ENV_ASMC_SIMPLE, "usv-asmc-v0": nat.ENV_ASMC_V0, "usv-simple": nat.ENV_SIMPLE,
             "usv-asmc-ye-int-v0": nat.ENV_ASMC_YE_INT_V0, "usv-pid-v0": nat.ENV_PID_V0,
             "usv-asmc-ca-v0": nat.ENV_ASMC_CA_V0}
LEGACY_IDS = ("usv-asmc-v0", "usv-asmc-ye-int-v0", "usv-pid-v0")  # UsvAsmcEnv's step() and its two float64 siblings


class LazyInfos(Sequence):
    """infos of one step without building N dicts -- or any dict before it is asked for: only the envs that finished
    carry content (SB3's ``episode`` / ``TimeLimit.truncated`` / ``terminal_observation``), kept as arrays over the done
    envs and turned into a dict on access."""

    def __init__(self, n, done_idx, time_limit, ep_return, ep_length, terminal_obs=None):
        self._n, self._idx = n, np.asarray(done_idx, dtype=np.int64)  # ascending (np.flatnonzero)
        self._tl, self._er, self._el, self._tob = time_limit, ep_return, ep_length, terminal_obs
        self._extra = {}

    def __len__(self):
        return self._n

    def _payload(self, k):
        d = {"TimeLimit.truncated": bool(self._tl[k]), "episode": {"r": float(self._er[k]), "l": int(self._el[k])}}
        if self._tob is not None:
            d["terminal_observation"] = self._tob[k]
        d.update(self._extra.get(k, {}))
        return d

    def set_extra(self, k, key, value):
        """attach ``key`` to the k-th done env's dict (used by the host-side resets of a custom experiment)"""
        self._extra.setdefault(k, {})[key] = value

    def done_items(self):
        """(env index, info dict) of the envs that finished in the step"""
        return [(int(i), self._payload(k)) for k, i in enumerate(self._idx)]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(self._n))]
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError(i)
        k = int(np.searchsorted(self._idx, i))
        if k < self._idx.size and self._idx[k] == i:
            return self._payload(k)
        return {}


class UsvGpuVecEnv:
    metadata = {"render_modes": ["human", "rgb_array"], "render_fps": 30}  # simple_env.py:8

    def __init__(self, num_envs: int, device: int | str | torch.device = 0, seed: int = 0, *,
                 ignore_obstacles: bool = False, do_perturb: bool = False, skip_kinematic: bool = False,
                 max_episode_steps: Optional[int] = None, auto_reset: bool = True, precision: str = "f64",
                 obstacles_f32: bool = False, global_env_offset: int = 0, info_mode: str = "list",
                 render_mode: Optional[str] = None, copy_host_outputs: bool = False, mapped_obs: bool = False,
                 host_chunks: int = 2, env_id: str = ENV_ID, options: Optional[dict] = None, host_obs: bool = True):
        if not torch.cuda.is_available():
            raise nat.UsvError("UsvGpuVecEnv needs a CUDA device: there is no CPU fallback")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        if self.device.type != "cuda":
            raise nat.UsvError("UsvGpuVecEnv needs a CUDA device: there is no CPU fallback")
        self.num_envs = int(num_envs)
        if env_id not in ENV_KINDS:
            raise nat.UsvError(f"unknown env id '{env_id}' (have: {sorted(ENV_KINDS)})")
        self.env_id = env_id
        self.legacy = env_id in LEGACY_IDS  # UsvAsmcEnv, gym_usv/envs/usv_asmc_env.py, UsvAsmcYeIntEnv, UsvPidEnv
        self.legacy64 = self.legacy and env_id != "usv-asmc-v0"  # float64 state (USV_FIELD_LEGACY64)
        self.ca = env_id == "usv-asmc-ca-v0"  # UsvAsmcCaEnv, gym_usv/envs/usv_asmc_ca_env.py (in-repo UsvAsmc substituted)
        self.obs_dim = nat.LEGACY_OBS_DIM if self.legacy else nat.CA_OBS_DIM if self.ca else nat.OBS_DIM
        self.act_dim = 1 if self.legacy else nat.ACT_DIM
        self.observation_space = (spaces.legacy_observation_space() if self.legacy else spaces.ca_observation_space()
                                  if self.ca else spaces.simple_observation_space())
        self.action_space = (spaces.legacy_action_space() if self.legacy else spaces.ca_action_space() if self.ca
                             else spaces.simple_action_space())
        self.render_mode = render_mode
        self.render_envs = 16   # how many envs get_images() / render() draw
        self.render_size = 512  # SimpleEnvVisualizer(window_size=512), simple_env_visualizer.py:9
        self.info_mode = info_mode
        # UsvSimpleEnv.__init__(render_mode, options) (simple_env.py:10,15): 'run_custom_experiment' + 'experiment' make
        # EVERY reset of that env install a fixed scene (:292-300).  Only 'usv-simple' takes constructor options
        # (UsvSimpleASMCEnv.__init__ has no such argument, simple_env_asmc.py:10-11).  With an experiment installed the
        # resets of finished envs run through the host (in-kernel auto-reset off), see step_wait().
        self.options = dict(options or {})
        self._experiment = None
        if self.options.get("run_custom_experiment"):
            if env_id != "usv-simple":
                raise nat.UsvError("constructor options (run_custom_experiment) exist for 'usv-simple' only "
                                   "(gym_usv/envs/simple_env.py:10; simple_env_asmc.py:10-11 takes none)")
            self._experiment = self.options["experiment"]
        self._host_auto_reset = bool(auto_reset) and self._experiment is not None
        if self._host_auto_reset:
            auto_reset = False
        if max_episode_steps is None:
            max_episode_steps = MAX_EPISODE_STEPS[env_id]
        if precision != "f64":
            # USV_PREC_MIXED is a reserved value of the C ABI, not an implemented mode: the parity bar (exact done flags
            # against the reference's float64 arithmetic) does not survive float32 force terms
            raise nat.UsvError(f"precision '{precision}' is not implemented: this build computes in float64 ('f64') only")
        self.cfg = nat.UsvConfig(ENV_KINDS[env_id], int(ignore_obstacles), int(do_perturb), int(skip_kinematic),
                                 int(max_episode_steps), int(auto_reset), nat.PREC_F64, int(obstacles_f32))
        self._seed = int(seed)
        self.global_env_offset = int(global_env_offset)
        self._lib = nat.lib()
        self._h = C.c_void_p()
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self._dev_index = dev_index
        nat.check(self._lib.usvgpu_create(C.byref(self.cfg), dev_index, self.num_envs, self.global_env_offset,
                                          self._seed, C.byref(self._h)))
        n = self.num_envs
        with torch.cuda.device(self.device):
            self.obs = torch.zeros((n, self.obs_dim), dtype=torch.float32, device=self.device)
            self.terminal_obs = torch.zeros((n, self.obs_dim), dtype=torch.float32, device=self.device)
            self.reward = torch.zeros(n, dtype=torch.float32, device=self.device)
            self._flags = torch.zeros((2, n), dtype=torch.uint8, device=self.device)
            self.terminated, self.truncated = self._flags[0], self._flags[1]
            self._actions_dev = torch.zeros((n, self.act_dim), dtype=torch.float32, device=self.device)
        # page-locked host buffers of the numpy face, allocated on first use: two sets, used alternately, so the
        # arrays step() returned stay valid while the next step is computed (copy_host_outputs=True hands out
        # private copies instead)
        self._host = None
        self._slot = 0
        self.copy_host_outputs = bool(copy_host_outputs)
        self.mapped_obs = bool(mapped_obs)
        # host_obs=False: step() / step_wait() return the observations as the env's CUDA tensor (self.obs) instead of
        # a numpy array -- reward / dones / infos still come back on the host.  For a policy that lives on the GPU
        # under a host-side training loop: 6 bytes per env-step cross PCIe instead of 578.
        self.host_obs = bool(host_obs)
        if not self.host_obs and self._host_auto_reset:
            raise nat.UsvError("host_obs=False cannot be combined with a custom experiment: its host-side resets read the "
                               "observation rows on the host")
        self.host_chunks = int(host_chunks)
        self._bind_staging()
        self._pending = None
        self._closed = False

    # ------------------------------------------------------------------ plumbing
    def _bind_staging(self):
        self._check(self._lib.usvgpu_set_host_chunks(self._h, self.host_chunks))
        # the host-buffer face stages through the same device tensors as the tensor face
        nat.check(self._lib.usvgpu_use_staging(self._h, *[C.c_void_p(t.data_ptr()) for t in (
            self._actions_dev, self.obs, self.reward, self._flags, self.terminal_obs)]), self._h)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _check(self, rc):
        nat.check(rc, self._h)

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def close(self):
        if not self._closed and self._h:
            if self._host is not None:
                torch.cuda.synchronize(self.device)
                for ptr in self._host_ptrs:
                    self._lib.usvgpu_host_free(ptr)
                self._host = None
            self._lib.usvgpu_destroy(self._h)
            self._closed = True

    def _host_array(self, shape, dtype):
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = C.c_void_p()
        nat.check(self._lib.usvgpu_host_alloc(C.byref(ptr), nbytes))
        self._host_ptrs.append(ptr)
        buf = (C.c_char * nbytes).from_address(ptr.value)
        arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
        arr[...] = 0
        return arr

    def _host_buffers(self):
        if self._host is None:
            n = self.num_envs
            self._host_ptrs = []
            self._host = [{"actions": self._host_array((n, self.act_dim), np.float32),
                           "obs": self._host_array((n, self.obs_dim), np.float32),
                           "reward": self._host_array((n,), np.float32),
                           "term": self._host_array((n,), np.uint8),
                           "trunc": self._host_array((n,), np.uint8)} for _ in range(2)]
        return self._host

    # ------------------------------------------------------------------ device-resident face
    def reset_tensor(self, mask: Optional[torch.Tensor] = None, options: Optional[dict] = None) -> torch.Tensor:
        """Env.reset() for all envs (or those with mask != 0).  Returns the device obs tensor [N,143]
        (columns 15.. keep their previous content: simple_env.py:302 reuses stale sensor data).  ``options``: for
        'usv-asmc-ca-v0' the option dict of UsvAsmcCaEnv.reset (usv_asmc_ca_env.py:358-372), applied in the kernel."""
        mptr = None
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            mptr = C.c_void_p(mask.data_ptr())
        if options and self.ca:
            opt = nat.ca_reset_options(options)
            self._check(self._lib.usvgpu_reset_ca_options(self._h, mptr, C.byref(opt), C.c_void_p(self.obs.data_ptr()),
                                                          self._stream()))
            return self.obs
        self._check(self._lib.usvgpu_reset(self._h, mptr, C.c_void_p(self.obs.data_ptr()), self._stream()))
        return self.obs

    def step_tensor(self, actions: torch.Tensor):
        """actions: float32 CUDA tensor [N,2].  Returns (obs, reward, terminated, truncated) device
        tensors owned by the env (overwritten by the next step).  No host copy, no sync."""
        if actions.device != self.device or actions.dtype != torch.float32 or not actions.is_contiguous():
            actions = actions.to(device=self.device, dtype=torch.float32).contiguous()
        if self.legacy and actions.dim() == 1:
            actions = actions.unsqueeze(1)
        if actions.shape != (self.num_envs, self.act_dim):
            raise ValueError(f"actions must have shape {(self.num_envs, self.act_dim)}, got {tuple(actions.shape)}")
        self._check(self._lib.usvgpu_step(
            self._h, C.c_void_p(actions.data_ptr()), C.c_void_p(self.obs.data_ptr()),
            C.c_void_p(self.reward.data_ptr()), C.c_void_p(self.terminated.data_ptr()),
            C.c_void_p(self.truncated.data_ptr()), C.c_void_p(self.terminal_obs.data_ptr()), self._stream()))
        return self.obs, self.reward, self.terminated, self.truncated

    def step_tensor_timed(self, actions: torch.Tensor):
        """step_tensor() that synchronises and returns (dynamics_kernel_ms, sensor_kernel_ms) measured with CUDA
        events on the launching stream (bench.py roofline line)."""
        assert actions.is_cuda and actions.dtype == torch.float32 and actions.is_contiguous()
        ms = (C.c_float * 2)()
        self._check(self._lib.usvgpu_step_timed(
            self._h, C.c_void_p(actions.data_ptr()), C.c_void_p(self.obs.data_ptr()),
            C.c_void_p(self.reward.data_ptr()), C.c_void_p(self.terminated.data_ptr()),
            C.c_void_p(self.truncated.data_ptr()), C.c_void_p(self.terminal_obs.data_ptr()), self._stream(), ms))
        return float(ms[0]), float(ms[1])

    def rollout_tensor(self, actions: torch.Tensor, n_steps: int, action_wrap: int = 0):
        """n_steps launches with pre-staged actions [K,N,2] (tools/profile_env.py:5-7 style loop)."""
        assert actions.is_cuda and actions.dtype == torch.float32 and actions.is_contiguous()
        self._check(self._lib.usvgpu_rollout(
            self._h, C.c_void_p(actions.data_ptr()), int(n_steps), int(action_wrap), C.c_void_p(self.obs.data_ptr()),
            C.c_void_p(self.reward.data_ptr()), C.c_void_p(self.terminated.data_ptr()),
            C.c_void_p(self.truncated.data_ptr()), self._stream()))
        return self.obs, self.reward, self.terminated, self.truncated

    # ------------------------------------------------------------------ SB3 VecEnv face (host numpy)
    def reset(self, seed: Optional[int] = None, options: Optional[dict] = None) -> np.ndarray:
        """VecEnv.reset().  ``options`` are the reset options of the reference env, applied to every env: for
        'usv-simple' UsvSimpleEnv.reset(seed, options) (simple_env.py:228,276-288: ``{'place_obstacles_on_path': n}``),
        for 'usv-asmc-ca-v0' UsvAsmcCaEnv.reset(options=...) (usv_asmc_ca_env.py:358-372: obs_x / obs_y / obs_r,
        target_point, start_position); 'usv-asmc-simple' drops them like the reference does (simple_env_asmc.py:14-16
        passes only the seed on), the legacy envs take none."""
        if seed is not None:
            self.seed(seed)
        if options is None and getattr(self, "_next_reset_options", None):
            options, self._next_reset_options = self._next_reset_options, None
        hb = self._host_buffers()[self._slot]
        if self.ca and options:  # usv_asmc_ca_env.py:358-372, applied inside the reset kernel
            self.reset_tensor(options=options)
            hb["obs"][...] = self.obs.cpu().numpy()
            return hb["obs"].copy() if self.copy_host_outputs else hb["obs"]
        if self.env_id != "usv-simple":
            options = None
        if options or self._experiment is not None:
            self.reset_tensor()
            self._apply_reset_edits(None, options)
            hb["obs"][...] = self.obs.cpu().numpy()
        else:
            self._check(self._lib.usvgpu_reset_host(self._h, None, hb["obs"].ctypes.data_as(C.c_void_p), self._stream()))
        return hb["obs"].copy() if self.copy_host_outputs else hb["obs"]

    def set_options(self, options=None) -> None:
        """SB3 VecEnv.set_options(): options for the next reset() (one dict for all envs, or a list whose first entry is
        used: the batched env resets all envs alike)."""
        if isinstance(options, (list, tuple)):
            options = options[0] if options else None
        self._next_reset_options = dict(options) if options else None

    # ------------------------------------------------------------------ reset options / custom experiment (host edits)
    def _apply_reset_edits(self, idx, options: Optional[dict]) -> None:
        """The reference's option-driven parts of reset(), as edits of a fresh in-kernel reset (simple_env.py:276-300):
        obstacles placed along the path ('place_obstacles_on_path': n, drawn here with numpy keyed by (seed, global env
        id, episode)) and the constructor's custom experiment (fixed obstacles, path, pose).  Then the reset observation
        of the edited envs is recomputed on the device (usvgpu_reset_obs)."""
        idx = np.arange(self.num_envs) if idx is None else np.asarray(idx, dtype=np.int64)
        if idx.size == 0:
            return
        st = self.get_state(["dyn", "epi", "obstacles", "n_obs", "episode"])
        if options and options.get("place_obstacles_on_path"):
            n_add = int(options["place_obstacles_on_path"])
            for i in idx:
                k0 = int(st["n_obs"][i])
                if k0 + n_add > nat.MAX_OBSTACLES:
                    raise nat.UsvError(f"place_obstacles_on_path={n_add}: env {i} would hold {k0 + n_add} obstacles, "
                                       f"the table has {nat.MAX_OBSTACLES} slots")
                rng = np.random.default_rng([self._seed & 0xFFFFFFFF, self.global_env_offset + int(i),
                                             int(st["episode"][i]) & 0xFFFFFFFF, 0x0B57])
                ps, pe = st["epi"][0:2, i], st["epi"][2:4, i]
                angle = np.arctan2(pe[1] - ps[1], pe[0] - ps[0])
                mag = rng.uniform(0, np.hypot(0.0, 20.0), n_add)                       # :281
                st["obstacles"][i, 0, k0:k0 + n_add] = rng.normal(np.cos(angle) * mag + ps[0], 1)   # :283
                st["obstacles"][i, 1, k0:k0 + n_add] = rng.normal(np.sin(angle) * mag + ps[1], 1)   # :284
                st["obstacles"][i, 2, k0:k0 + n_add] = rng.uniform(0.15, 0.5, n_add)                 # :290
                st["n_obs"][i] = k0 + n_add
        if self._experiment is not None:
            exp = self._experiment
            pos = np.asarray(exp["obstacle_positions"], dtype=np.float64).reshape(-1, 2)
            rad = np.asarray(exp["obstacle_radius"], dtype=np.float64).reshape(-1)
            if pos.shape[0] > nat.MAX_OBSTACLES or rad.shape[0] != pos.shape[0]:
                raise nat.UsvError(f"custom experiment: {pos.shape[0]} obstacles / {rad.shape[0]} radii; the table has "
                                   f"{nat.MAX_OBSTACLES} slots")
            ps = np.asarray(exp["path_start"], dtype=np.float64).reshape(2)
            pe = ps + np.array([np.cos(exp["angle"]), np.sin(exp["angle"])]) * 100                   # :297
            p0 = np.asarray(exp["position"], dtype=np.float64).reshape(3)
            k = pos.shape[0]
            for i in idx:
                st["obstacles"][i, 0, :k], st["obstacles"][i, 1, :k], st["obstacles"][i, 2, :k] = pos[:, 0], pos[:, 1], rad
                st["n_obs"][i] = k
                st["epi"][0:2, i], st["epi"][2:4, i] = ps, pe
                st["dyn"][0:3, i] = p0
        self.set_state({k: st[k] for k in ("dyn", "epi", "obstacles", "n_obs")})
        mask = torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device)
        mask[torch.as_tensor(idx, device=self.device)] = 1
        self._check(self._lib.usvgpu_reset_obs(self._h, C.c_void_p(mask.data_ptr()), C.c_void_p(self.obs.data_ptr()),
                                               self._stream()))
        torch.cuda.current_stream(self.device).synchronize()

    def refresh_reset_obs(self, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """obs[:, :15] of the envs in ``mask`` (default: all) recomputed as the reset observation of their current
        state (usvgpu_reset_obs): for a caller that edits a fresh reset through set_state()."""
        mptr = None
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            mptr = C.c_void_p(mask.data_ptr())
        self._check(self._lib.usvgpu_reset_obs(self._h, mptr, C.c_void_p(self.obs.data_ptr()), self._stream()))
        return self.obs

    # ------------------------------------------------------------------ per-step info (simple_env.py:102-115,189-199)
    def step_info_tensor(self, indices=None) -> torch.Tensor:
        """float64 CUDA tensor [K, 24]: the reference's info dict as columns (include/usvgpu.h UsvInfoCol), recomputed by
        usvgpu_step_info from the current state of the envs in ``indices`` (default: all)."""
        if indices is None:
            idx, k = None, self.num_envs
        else:
            idx = torch.as_tensor(indices, dtype=torch.int64, device=self.device).reshape(-1).contiguous()
            k = int(idx.numel())
        out = torch.empty((k, nat.INFO_DIM), dtype=torch.float64, device=self.device)
        if k:
            self._check(self._lib.usvgpu_step_info(self._h, C.c_void_p(idx.data_ptr()) if idx is not None else None, k,
                                                   C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def step_info(self, indices=None) -> List[dict]:
        """The info dicts UsvSimpleEnv.step returns (``info = _get_info(reward, action); info.update(reward_info)``,
        simple_env.py:341-342) for the envs in ``indices``, materialised on request from the device matrix."""
        rows = self.step_info_tensor(indices).cpu().numpy()
        sel = np.arange(self.num_envs) if indices is None else np.asarray(indices).reshape(-1)
        rew = self.reward[torch.as_tensor(sel, device=self.device)].cpu().numpy() if rows.shape[0] else []
        out = []
        for k in range(rows.shape[0]):
            d = {name: (rows[k, c0:c0 + w].copy() if w > 1 else float(rows[k, c0])) for name, (c0, w) in nat.INFO_KEYS.items()}
            d.update(reward=float(rew[k]), left_thruster=0, right_thruster=0)
            out.append(d)
        return out

    def step_async(self, actions) -> None:
        self._slot ^= 1
        hb = self._host_buffers()[self._slot]
        hb["actions"][...] = np.asarray(actions, dtype=np.float32).reshape(self.num_envs, self.act_dim)
        p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        self._check(self._lib.usvgpu_step_host(self._h, p(hb["actions"]), p(hb["obs"]) if self.host_obs else None,
                                               p(hb["reward"]), p(hb["term"]), p(hb["trunc"]),
                                               int(self.mapped_obs and self.host_obs), self._stream()))
        self._pending = True

    def step_wait(self):
        assert self._pending, "step_wait() without step_async()"
        # reward / flags arrive first: the episode bookkeeping below overlaps the observation copy
        self._check(self._lib.usvgpu_flags_sync(self._h))
        self._pending = None
        hb = self._host[self._slot]
        rew = hb["reward"].copy() if self.copy_host_outputs else hb["reward"]
        term = hb["term"].view(np.bool_)
        trunc = hb["trunc"].view(np.bool_)
        dones = term | trunc
        done_idx = np.flatnonzero(dones)
        tob = None
        tl = trunc[done_idx] & ~term[done_idx]
        epi = np.empty((done_idx.size, 2), dtype=np.float64)
        if done_idx.size:
            # terminal observations and the (return, length) of the finished episodes, both kept on the device
            idx32 = done_idx.astype(np.int32)
            tob = np.empty((done_idx.size, self.obs_dim), dtype=np.float32)
            self._check(self._lib.usvgpu_terminal_obs_host(self._h, idx32.ctypes.data_as(C.c_void_p), idx32.size,
                                                           tob.ctypes.data_as(C.c_void_p),
                                                           epi.ctypes.data_as(C.c_void_p)))
            if not self.cfg.auto_reset and not self._host_auto_reset:
                tob = None  # nothing was reset: SB3's key would be misleading
        lazy = LazyInfos(self.num_envs, done_idx, tl, epi[:, 0], epi[:, 1], tob)
        self._check(self._lib.usvgpu_stream_sync(self._h, self._stream()))  # the observations are on the host
        if self._host_auto_reset and done_idx.size:
            # a custom experiment is installed (constructor options): the finished envs are reset here, through the
            # same edits as reset(), and get SB3's terminal_observation from the rows the step just wrote
            lazy._tob = hb["obs"][done_idx].copy()
            mask = torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device)
            mask[torch.as_tensor(done_idx, device=self.device)] = 1
            self.reset_tensor(mask)
            self._apply_reset_edits(done_idx, None)
            hb["obs"][done_idx] = self.obs[torch.as_tensor(done_idx, device=self.device)].cpu().numpy()
        obs = hb["obs"].copy() if self.copy_host_outputs else hb["obs"]
        if not self.host_obs:
            obs = self.obs  # device-resident rows (the staging tensor of the host face is this very tensor)
        if self.info_mode == "lazy":
            infos: Any = lazy
        else:
            infos = [{} for _ in range(self.num_envs)]
            for i, p in lazy.done_items():
                infos[i] = p
        return obs, rew, dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    # ------------------------------------------------------------------ VecEnv odds and ends
    def seed(self, seed: Optional[int] = None) -> List[Optional[int]]:
        """Re-keys the reset RNG: takes effect for envs reset from now on (new handle state is kept)."""
        if seed is not None:
            self._seed = int(seed)
            # the key lives in the handle: recreate it, keeping nothing (SB3 calls seed() before reset())
            self._lib.usvgpu_destroy(self._h)
            self._h = C.c_void_p()
            nat.check(self._lib.usvgpu_create(C.byref(self.cfg), self._dev_index, self.num_envs,
                                              self.global_env_offset, self._seed, C.byref(self._h)))
            self._bind_staging()
        return [None if seed is None else seed + i for i in range(self.num_envs)]

    def _indices(self, indices) -> Iterable[int]:
        if indices is None:
            return range(self.num_envs)
        if isinstance(indices, int):
            return [indices]
        return indices

    def get_attr(self, attr_name: str, indices=None) -> List[Any]:
        idx = list(self._indices(indices))
        if attr_name in ("render_mode", "observation_space", "action_space", "metadata"):
            return [getattr(self, attr_name)] * len(idx)
        fields = self.get_state()
        if attr_name in ("position", "velocity"):
            if self.legacy:  # rows of USV_FIELD_LEGACY / LEGACY64: action_last | u v r | x y psi | ...
                st = fields["legacy64" if self.legacy64 else "legacy"]
                rows = st[4:7] if attr_name == "position" else st[1:4]
            else:
                rows = fields["dyn"][0:3] if attr_name == "position" else fields["dyn"][3:6]
            return [rows[:, i].copy() for i in idx]
        raise AttributeError(attr_name)

    def set_attr(self, attr_name: str, value: Any, indices=None) -> None:
        raise AttributeError(f"cannot set '{attr_name}' on a GPU env; use set_state()")

    def env_method(self, method_name: str, *args, indices=None, **kwargs) -> List[Any]:
        raise AttributeError(f"'{method_name}' is not available on a GPU env")

    def env_is_wrapped(self, wrapper_class, indices=None) -> List[bool]:
        return [False] * len(list(self._indices(indices)))

    # ------------------------------------------------------------------ rgb_array frames
    def render_tensor(self, indices=None, size: Optional[int] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """UsvSimpleEnv.render() with render_mode "rgb_array" (simple_env.py:117-131 -> simple_env_visualizer.py:17-114)
        for the envs in `indices` (default: the first ``render_envs``): uint8 CUDA tensor [K, size, size, 3] drawn by
        usvgpu_render from the device state and the last observation.  No window system involved."""
        if self.legacy or self.ca:
            raise NotImplementedError(f"{self.env_id} has no SimpleEnvVisualizer; its own viewer is unavailable on this box")
        size = int(size or self.render_size)
        if indices is None:
            idx, k = None, min(self.num_envs, self.render_envs)
        else:
            idx = torch.as_tensor(indices, dtype=torch.int64, device=self.device).reshape(-1).contiguous()
            k = int(idx.numel())
            if k and (int(idx.min()) < 0 or int(idx.max()) >= self.num_envs):
                raise IndexError("render_tensor: env index out of range")
        if out is None:
            out = torch.empty((k, size, size, 3), dtype=torch.uint8, device=self.device)
        assert out.is_contiguous() and out.dtype == torch.uint8 and tuple(out.shape) == (k, size, size, 3)
        self._check(self._lib.usvgpu_render(self._h, C.c_void_p(idx.data_ptr()) if idx is not None else None, k,
                                            C.c_void_p(self.obs.data_ptr()), int(size), C.c_void_p(out.data_ptr()),
                                            self._stream()))
        return out

    def get_images(self):
        """SB3 VecEnv.get_images(): frames of the first ``render_envs`` envs (tiling all N of a GPU-sized batch into
        one picture is of no use; set ``render_envs`` to change it)."""
        return list(self.render_tensor().cpu().numpy())

    def render(self, mode: Optional[str] = None):
        """SB3 VecEnv.render(): the get_images() frames tiled into one rgb_array (tile_images layout: as square as
        possible).  "human" mode needs a window system: unavailable on this box."""
        if (mode or self.render_mode) != "rgb_array":
            return None
        imgs = self.render_tensor()
        k, hh, ww, _ = imgs.shape
        rows = int(np.ceil(np.sqrt(k)))
        cols = int(np.ceil(k / rows))
        canvas = torch.zeros((rows * cols, hh, ww, 3), dtype=torch.uint8, device=self.device)
        canvas[:k] = imgs
        return canvas.reshape(rows, cols, hh, ww, 3).permute(0, 2, 1, 3, 4).reshape(rows * hh, cols * ww, 3).cpu().numpy()

    # ------------------------------------------------------------------ state access (checkpoint / parity)
    # name -> (field id, dtype, leading dims before N, trailing dims after N)
    _FIELDS = {"dyn": (nat.FIELD_DYN, np.float64, (nat.DYN_ROWS,), ()), "epi": (nat.FIELD_EPI, np.float64, (nat.EPI_ROWS,), ()),
               "obstacles": (nat.FIELD_OBSTACLES, None, (), (3, nat.MAX_OBSTACLES)),
               "n_obs": (nat.FIELD_N_OBS, np.int32, (), ()), "target0": (nat.FIELD_TARGET0, np.float32, (), (2,)),
               "steps": (nat.FIELD_STEPS, np.int32, (), ()),
               "episode": (nat.FIELD_EPISODE, np.int32, (), ()), "ep_return": (nat.FIELD_EP_RETURN, np.float64, (), ())}
    _LEGACY_FIELDS = {"legacy": (nat.FIELD_LEGACY, np.float32, (nat.LEGACY_ROWS,), ()),
                      "target": (nat.FIELD_LEGACY_TARGET, np.float64, (6,), ()),
                      "pos0": (nat.FIELD_LEGACY_POS0, np.float64, (3,), ()),
                      "steps": (nat.FIELD_STEPS, np.int32, (), ()), "episode": (nat.FIELD_EPISODE, np.int32, (), ()),
                      "ep_return": (nat.FIELD_EP_RETURN, np.float64, (), ())}
    _LEGACY64_FIELDS = {"legacy64": (nat.FIELD_LEGACY64, np.float64, (nat.LEGACY64_ROWS,), ()),
                        **{k: v for k, v in _LEGACY_FIELDS.items() if k not in ("legacy", "pos0")}}

    _CA_FIELDS = {"dyn": _FIELDS["dyn"], "ca": (nat.FIELD_CA, np.float64, (nat.CA_ROWS,), ()),
                  "n_obs": (nat.FIELD_CA_N_OBS, np.int32, (), ()),
                  "steps": _FIELDS["steps"], "episode": _FIELDS["episode"], "ep_return": _FIELDS["ep_return"]}

    def _field_table(self):
        if self.ca:
            return self._CA_FIELDS
        return self._LEGACY64_FIELDS if self.legacy64 else self._LEGACY_FIELDS if self.legacy else self._FIELDS

    def _field_array(self, name):
        fid, dt, lead, trail = self._field_table()[name]
        if dt is None:
            dt = np.float32 if self.cfg.obstacles_f32 else np.float64
        return fid, np.zeros(lead + (self.num_envs,) + trail, dtype=dt)

    def get_state(self, names: Optional[Sequence[str]] = None) -> dict:
        """Host copies of the state field groups ([rows][N]; obstacles [N][3][32]); see include/usvgpu.h UsvField."""
        out = {}
        for name in (names or self._field_table()):
            fid, arr = self._field_array(name)
            self._check(self._lib.usvgpu_get_state(self._h, fid, arr.ctypes.data_as(C.c_void_p), arr.nbytes,
                                                   self._stream()))
            out[name] = arr
        return out

    def set_state(self, state: dict) -> None:
        for name, val in state.items():
            fid, arr = self._field_array(name)
            arr[...] = val
            self._check(self._lib.usvgpu_set_state(self._h, fid, arr.ctypes.data_as(C.c_void_p), arr.nbytes,
                                                   self._stream()))

    def episode_stats(self, reset: bool = False) -> dict:
        st = nat.UsvStats()
        self._check(self._lib.usvgpu_episode_stats(self._h, C.byref(st), int(reset), self._stream()))
        return st.as_dict()

    def stats_tensor(self) -> torch.Tensor:
        """The device-side UsvStats block as 8 raw 64-bit words (int64 view, no copy): words 0-1 are the float64
        return / length sums (``.view(torch.float64)[:2]``), words 2-7 the counters episodes, terminated,
        out_of_bounds, time_limit, env_steps (kept on the host: 0 here), nonfinite.  For a caller that wants to
        all-reduce the statistics with NCCL without a host round trip."""
        ptr = C.c_void_p()
        self._check(self._lib.usvgpu_stats_device_ptr(self._h, C.byref(ptr)))

        class _Block:
            __cuda_array_interface__ = {"shape": (8,), "typestr": "<i8", "data": (int(ptr.value), False), "version": 2}

        return torch.as_tensor(_Block(), device=self.device)
