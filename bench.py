#!/usr/bin/env python
"""bench.py -- env-steps/s of the batched 'usv-asmc-simple' step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W]          # our CUDA path
    python bench.py --impl reference [...]                        # the reference's CPU algorithm (oracle port)
    torchrun --nproc-per-node N ... bench.py --gpus N ...         # one rank per GPU, envs sharded by index range

A "step" is one VecEnv.step() over one batch of envs on every rank (two kernel launches per rank:
dynamics_kernel = 20 ASMC + 3-DOF sub-steps, kinematic post-step, closest point, header, partial reward;
sensor_kernel = 128-ray sensor pass, flags, reward, in-kernel auto-reset, observation rows).  Workload: BASELINE.json configs[2] -- 65 536 envs per GPU with randomised
obstacles and range-sensor observations (weak scaling: 8 GPUs = 524 288 envs; --envs-per-gpu 131072 gives
the 1 M-env configs[3]).  The timed loop rotates over --batches independent env batches so that the state a
step reads was evicted from the 126 MB L2 by the other batches' steps (working set > L2).

Printed JSON (one line, rank 0): see the keys below.  `value` = whole-job env-steps/s with actions and state
resident in HBM; `e2e` = the same metric through the host-numpy VecEnv.step() face (pinned H2D of actions,
D2H of observations / rewards / flags inside the timed region); `roofline` = algorithmic bytes per launch /
CUDA-event launch time against the measured HBM copy bandwidth; `cpu_baseline` = the CPU oracle (C port of the
reference's numpy algorithm) timed on this box's host cores on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "env_steps_per_sec"
UNIT = "env-steps/s"
# Algorithmic HBM bytes per env-step (DESIGN.md "Algorithmic bytes"; SURVEY.md 8d with float64 dynamic state):
#   read : 23 f64 dynamic state + 7 f64 episode constants + n_obs/steps (2 i32) + f64 episode return
#          + 32 x 3 obstacle table + 2 f32 action
#   write: 23 f64 dynamic state + steps i32 + f64 episode return + 143 f32 observation + f32 reward + 2 flag bytes
def algorithmic_bytes(obstacles: bool, obstacles_f32: bool) -> int:
    rd = 23 * 8 + 7 * 8 + 2 * 4 + 8 + 2 * 4
    if obstacles:
        rd += 32 * 3 * (4 if obstacles_f32 else 8)
    wr = 23 * 8 + 4 + 8 + 143 * 4 + 4 + 2
    return rd + wr


# Per-kernel share of those bytes plus the 61 B hand-over block each side moves (cos/sin of the heading 16,
# partial reward 8, 9 float32 header columns 36, flag byte 1): DESIGN.md "Kernels".
HANDOVER = 16 + 8 + 36 + 1


def kernel_bytes(obstacles: bool, obstacles_f32: bool, kinematic: bool = False) -> dict:
    dyn = (23 * 8 + 7 * 8 + 4 + 2 * 4) + (23 * 8 + 4) + HANDOVER
    if kinematic:  # 'usv-simple': pose, velocity, last_action (8 columns) + progress move; no controller memory
        dyn = (9 * 8 + 7 * 8 + 4 + 2 * 4) + (9 * 8 + 4) + HANDOVER
    sens = (2 * 8 + 4 + 8 + HANDOVER) + (143 * 4 + 4 + 2 + 8)
    if obstacles:
        sens += 32 * 3 * (4 if obstacles_f32 else 8)
    return {("usv::kinematic_kernel" if kinematic else "usv::dynamics_kernel"): dyn, "usv::sensor_kernel": sens}


# 'usv-asmc-v0' (UsvAsmcEnv, one sub-step per env-step): float32 SoA state [19], float64 target (4 of 6 rows read)
#   read : 19 f32 state + 4 f64 target + f32 action + i32 steps + f64 episode return            = 124 B
#   write: 19 f32 state + 6 f32 observation + f32 reward + 2 flag bytes + i32 steps + f64 return = 118 B
# SURVEY.md 8d "Algorithmic bytes per env-step" with float64 dynamic state (the storage this build uses)
SURVEY_BYTES_F64_STATE = {"obstacles": 1390, "no_obstacles": 1002}
LEGACY_BYTES = (19 * 4 + 4 * 8 + 4 + 4 + 8) + (19 * 4 + 6 * 4 + 4 + 2 + 4 + 8)
# the float64 siblings keep float64 state columns: 21 (ye-int) / 19 (pid) rows read and written
LEGACY_IDS = ("usv-asmc-v0", "usv-asmc-ye-int-v0", "usv-pid-v0")


def legacy_bytes(env_id: str) -> int:
    if env_id == "usv-asmc-v0":
        return LEGACY_BYTES
    rows = 21 if env_id == "usv-asmc-ye-int-v0" else 19
    return (rows * 8 + 4 * 8 + 4 + 4 + 8) + (rows * 8 + 6 * 4 + 4 + 2 + 4 + 8)


def legacy_kernel(env_id: str) -> str:
    return {"usv-asmc-v0": "usv::legacy_step_kernel", "usv-asmc-ye-int-v0": "usv::legacy64_step_kernel<1>",
            "usv-pid-v0": "usv::legacy64_step_kernel<2>"}[env_id]


def ncu_traffic(kernel: str, n_envs: int):
    """dram bytes per launch of `kernel` from the committed ncu --set full capture (profiles/ncu_traffic.json), or None
    when the capture was taken at another batch size."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            d = json.load(f)
        want = 1048576 if "legacy" in kernel else 65536
        return d.get(kernel) if n_envs == want else None
    except Exception:  # noqa: BLE001
        return None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU during the timed region (NVML, 20 ms period;
    nvidia-smi query as a fallback)."""

    def __init__(self, index: int, uuid: str | None = None):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = None
            if uuid:
                try:
                    self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
                except Exception:  # noqa: BLE001
                    self._h = None
            if self._h is None:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001
            self._nv = None

    _BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
             0x80: "hw_power_brake", 0x2: "app_clocks_setting", 0x10: "sync_boost", 0x100: "display_clocks"}

    def _sample_once(self):
        if self._nv is not None:
            nv = self._nv
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:  # noqa: BLE001
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for bit, name in self._BITS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            return
        try:
            out = subprocess.run(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm,"
                 "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                 "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.strip()
            f = [x.strip() for x in out.split(",")]
            self.samples.append(int(f[0]))
            self.max_mhz = int(f[1])
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:]):
                if v.lower().startswith("active"):
                    self.reasons.add(name)
        except Exception:  # noqa: BLE001
            pass

    def _run(self):
        while not self._stop.is_set():
            self._sample_once()
            self._stop.wait(0.02 if self._nv is not None else 0.2)

    def start(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def stop(self):
        self._sample_once()
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=5)
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
def cpu_oracle_rate(n_envs: int, budget_s: float, threads: int, ignore_obstacles: bool, seed: int = 0,
                    legacy: bool = False, env_id: str = "usv-asmc-v0"):
    """env-steps/s of the CPU oracle (all `threads` host threads) on n_envs lock-step envs, for ~budget_s."""
    import numpy as np
    import usv_oracle as uo
    rng = np.random.default_rng(0)
    if legacy:
        ov = uo.OracleLegacyVecEnv(n_envs, seed=seed, n_threads=threads, env_id=env_id)
        acts = rng.uniform(-np.pi / 2, np.pi / 2, (8, n_envs)).astype(np.float32)
    else:
        kin = env_id == "usv-simple"
        ov = uo.OracleVecEnv(n_envs, seed=seed, ignore_obstacles=ignore_obstacles, n_threads=threads, kinematic=kin,
                             max_episode_steps=500 if kin else 1000)
        acts = rng.uniform([0.2, -1], [1, 1], (8, n_envs, 2)).astype(np.float32)
    ov.reset()
    for k in range(2):
        ov.step(acts[k])
    t0 = time.perf_counter()
    steps = 0
    while True:
        ov.step(acts[steps % 8])
        steps += 1
        el = time.perf_counter() - t0
        if el >= budget_s:
            break
    return n_envs * steps / el, steps, el


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU algorithm for the path (the C oracle port of its numpy code; the
    reference itself is Python + gymnasium and cannot travel to the GPU box), all host threads, on a bounded
    sample of the same workload.  Rank 0 only."""
    if rank != 0:
        return
    import numpy as np
    import usv_oracle as uo
    threads = host_threads()
    n = args.ref_envs
    rng = np.random.default_rng(0)
    if args.env in LEGACY_IDS:
        ov = uo.OracleLegacyVecEnv(n, seed=args.seed, n_threads=threads, env_id=args.env)
        acts = rng.uniform(-np.pi / 2, np.pi / 2, (8, n)).astype(np.float32)
    else:
        kin = args.env == "usv-simple"
        ov = uo.OracleVecEnv(n, seed=args.seed, ignore_obstacles=bool(args.no_obstacles), n_threads=threads,
                             kinematic=kin, max_episode_steps=500 if kin else 1000)
        acts = rng.uniform([0.2, -1], [1, 1], (8, n, 2)).astype(np.float32)
    ov.reset()
    for k in range(args.warmup):
        ov.step(acts[k % 8])
    t0 = time.perf_counter()
    for k in range(args.steps):
        ov.step(acts[k % 8])
    el = time.perf_counter() - t0
    value = n * args.steps / el
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": dtype_of(args), "data": "synthetic",
        "config": workload_config(args, world, reference_sample=n),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{n} lock-step envs x {args.steps} steps, OpenMP over {threads} host threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def dtype_of(args):
    if args.env == "usv-asmc-v0":
        return "f32 state / f64 control terms (the reference's own mix)"
    if args.env in LEGACY_IDS:
        return "f64"
    return "f64"


def workload_config(args, world, reference_sample=None):
    if args.env in LEGACY_IDS:
        what = {"usv-asmc-v0": "legacy UsvAsmcEnv: ASMC path following", "usv-pid-v0": "UsvPidEnv: PID path following",
                "usv-asmc-ye-int-v0": "UsvAsmcYeIntEnv: ASMC path following with the cross-track integral"}[args.env]
        cfg = {"workload": f"{args.env} ({what} without obstacles, one 0.01 s sub-step "
                           f"per env-step), {args.envs_per_gpu} envs per GPU, BASELINE.json configs[1] secondary reading",
               "envs_per_gpu": args.envs_per_gpu, "envs_total": args.envs_per_gpu * world, "sub_steps_per_env_step": 1,
               "max_episode_steps": 0, "auto_reset": True,
               "sharding": f"contiguous env-index ranges over {world} GPU(s), no data-path collective",
               "actions": "uniform in the action box, pre-staged on device",
               "l2": f"{args.batches} rotating env batches per GPU"}
        if reference_sample is not None:
            cfg["reference_sample_envs"] = reference_sample
        return cfg
    kin = args.env == "usv-simple"
    cfg = {
        "workload": (("usv-simple (kinematic UsvSimpleEnv" if kin else "usv-asmc-simple (ASMC path following") +
                     (", obstacles ignored" if args.no_obstacles else
                     " with randomised obstacles + 128-ray range sensors") + f"), {args.envs_per_gpu} envs per GPU, " +
                     ("the env train_test/sb3_train_vec.py:55 trains on" if kin else "BASELINE.json configs[2]") +
                     ("" if args.envs_per_gpu == 65536 else " (resized)")),
        "envs_per_gpu": args.envs_per_gpu, "envs_total": args.envs_per_gpu * world,
        "sub_steps_per_env_step": 1 if kin else 20, "max_episode_steps": 500 if kin else 1000, "auto_reset": True,
        "precision": "f64", "obstacle_table": "f32" if args.obstacles_f32 else "f64",
        "sharding": f"contiguous env-index ranges over {world} GPU(s), no data-path collective",
        "actions": "uniform in the action box, pre-staged on device",
        "l2": f"{args.batches} rotating env batches per GPU (consecutive steps touch different state; working set > 126 MB L2)",
    }
    if reference_sample is not None:
        cfg["reference_sample_envs"] = reference_sample
    return cfg


# ------------------------------------------------------------------------------------------------
class Workload:
    """`batches` independent env shards of n envs on one GPU (rotated so that consecutive steps touch different state:
    working set > L2), pre-staged actions, and the K-step block in three forms: a Python loop of VecEnv.step_tensor()
    calls, the same loop captured once into a CUDA graph, and a timed replay of either."""

    def __init__(self, usv, torch, dev, local_rank, n, lo, args, env_id, no_obstacles, obstacles_f32, batches, seed):
        self.torch, self.dev, self.n = torch, dev, n
        self.envs = []
        for b in range(batches):
            e = usv.UsvGpuVecEnv(n, device=local_rank, seed=seed + b, global_env_offset=lo,
                                 ignore_obstacles=bool(no_obstacles), obstacles_f32=bool(obstacles_f32),
                                 info_mode="lazy", env_id=env_id)
            e.reset_tensor()
            self.envs.append(e)
        self.A = 8
        gen = torch.Generator(device=dev).manual_seed(1234 + lo)
        if env_id in LEGACY_IDS:
            self.acts = ((torch.rand((self.A, n, 1), generator=gen, device=dev) - 0.5) * 3.141592653589793).contiguous()
        else:
            a = torch.rand((self.A, n, 2), generator=gen, device=dev)
            self.acts = (a * torch.tensor([0.8, 2.0], device=dev) + torch.tensor([0.2, -1.0], device=dev)).contiguous()
        self.graph, self.graph_steps = None, 0

    def step(self, k):
        self.envs[k % len(self.envs)].step_tensor(self.acts[k % self.A])

    def loop(self, steps):
        for k in range(steps):
            self.step(k)

    def capture(self, steps):
        """The K-step block as ONE CUDA graph: the host enqueues it with a single call, so the timed region cannot be
        bound by the Python / ctypes launch loop.  Every step is still its own kernel launch(es) and the
        state round-trips HBM between steps -- nothing is fused across steps."""
        torch = self.torch
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self.loop(steps)
        self.graph, self.graph_steps = g, steps
        return g

    def preroll(self, steps_per_batch):
        """untimed steps so that episodes are spread over their whole life (steady-state mix of resets, collisions,
        out-of-bounds exits); replayed through a graph of one rotation to keep it short"""
        for k in range(len(self.envs)):  # one eager step per shard: nothing is loaded lazily inside a capture
            self.step(k)
        if steps_per_batch <= 0:
            return
        per = len(self.envs) * self.A  # one full rotation of shards x action slabs
        g = self.capture(per)
        for _ in range(max(1, (steps_per_batch * len(self.envs) + per - 1) // per)):
            g.replay()
        self.torch.cuda.synchronize(self.dev)

    def close(self):
        self.graph = None
        for e in self.envs:
            e.close()


def timed_blocks(torch, dist, dev, world, launch, repeats):
    """`repeats` times: barrier + synchronize, CUDA events around launch(), synchronize + barrier.  Returns per repeat the
    device ms (max over ranks) and this rank's host seconds spent inside launch() (enqueue time)."""
    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    ms_max, ms_own, host_s = [], [], []
    for _ in range(repeats):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        ev0.record()
        t0 = time.perf_counter()
        launch()
        host_s.append(time.perf_counter() - t0)
        ev1.record()
        sync_all()
        ms = ev0.elapsed_time(ev1)
        ms_own.append(ms)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max.append(float(t.item()))
    return ms_max, ms_own, host_s


def ncu_metrics():
    """What the committed ncu capture of the step kernel(s) says (profiles/ncu_metrics.json, written by
    tools/ncu_summary.py from the .ncu-rep of the same bench command): FP64 / issue-slot figures for the bench line."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_metrics.json")) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
        return {}


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    import gym_usv_b200 as usv
    from gym_usv_b200 import _native as nat
    from gym_usv_b200.sharding import reduce_episode_stats, shard_range

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # clocks are sampled from before the warm-up to after the last timed block (not started inside the timed window)
    try:
        uuid = "GPU-" + str(torch.cuda.get_device_properties(dev).uuid)
    except Exception:  # noqa: BLE001
        uuid = None
    sampler = ClockSampler(local_rank, uuid)
    sampler.start()

    n = args.envs_per_gpu
    lo, hi = shard_range(n * world, rank, world)
    legacy = args.env in LEGACY_IDS
    wl = Workload(usv, torch, dev, local_rank, hi - lo, lo, args, args.env, args.no_obstacles, args.obstacles_f32,
                  args.batches, args.seed)
    envs, acts, A = wl.envs, wl.acts, wl.A
    launches_per_step = 1 if legacy else 2

    wl.preroll(args.preroll)
    wl.loop(args.warmup)  # W untimed warm-up steps through the public per-step call
    torch.cuda.synchronize(dev)

    # ---- value: K steps, device-resident actions, CUDA events, max over ranks, median of R blocks -----------------
    K, R = args.steps, max(1, args.repeats)
    graph = wl.capture(K)
    graph.replay()  # one untimed replay (graph upload)
    ms_graph, ms_graph_own, host_graph = timed_blocks(torch, dist, dev, world, graph.replay, R)
    ms_loop, ms_loop_own, host_loop = timed_blocks(torch, dist, dev, world, lambda: wl.loop(K), R)
    ms_block = statistics.median(ms_graph)
    value = n * world * K / (ms_block * 1e-3)
    value_loop = n * world * K / (statistics.median(ms_loop) * 1e-3)

    # ---- e2e: host numpy in / host numpy out through VecEnv.step() -----------------------------------
    # usvgpu_step_host: pinned H2D of the actions, the kernels, D2H of observations / rewards / flags into
    # page-locked host buffers, stream sync, terminal observations of the finished envs -- all inside the timed region.
    e2e_steps = max(3, min(args.steps, args.e2e_steps))
    h_acts = np.ascontiguousarray(acts.cpu().numpy())

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def e2e_run(mapped):
        for e in envs:
            e.mapped_obs = bool(mapped)
        for k in range(3):
            envs[k % len(envs)].step(h_acts[k % A])
        sync_all()
        t0 = time.perf_counter()
        sink = 0.0
        for k in range(e2e_steps):
            obs, rew, dones, infos = envs[k % len(envs)].step(h_acts[k % A])
            sink += float(rew[0]) + float(obs[-1, -1])
        torch.cuda.synchronize(dev)
        te = time.perf_counter() - t0
        tt = torch.tensor([te], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return n * world * e2e_steps / float(tt.item())

    e2e_variants = {"staged_d2h": e2e_run(False)}
    if args.mapped_obs or args.e2e_all:
        e2e_variants["mapped_obs_stores"] = e2e_run(True)
    # not the headline: the same call with host_obs=False -- actions from the host, reward / flags / infos back to the
    # host, the observation rows left on the device for a GPU-resident policy (6 instead of 578 bytes per env-step D2H)
    for e in envs:
        e.host_obs = False
    e2e_variants["obs_left_on_device"] = e2e_run(False)
    for e in envs:
        e.host_obs = True
    for e in envs:
        e.mapped_obs = bool(args.mapped_obs)
    e2e_value = e2e_variants["mapped_obs_stores" if args.mapped_obs else "staged_d2h"]
    h2d = n * envs[0].act_dim * 4
    d2h = n * (envs[0].obs_dim * 4 + 4 + 1 + 1)

    # ---- per-kernel device time (CUDA events inside the library, on the launching stream) --------------
    kms = [0.0, 0.0]
    ksteps = max(3, min(args.steps, 60))
    for k in range(ksteps):
        d_ms, s_ms = envs[k % len(envs)].step_tensor_timed(acts[k % A])
        kms[0] += d_ms / ksteps
        kms[1] += s_ms / ksteps

    stats = envs[0].episode_stats()
    for e in envs[1:]:
        s = e.episode_stats()
        stats = {k: stats[k] + s[k] for k in stats}
    stats = reduce_episode_stats(stats, device=dev)
    wl.close()

    # ---- measured FP64 issue rate of this GPU (the binding resource of the float64 path) ----------------
    dfma = C_double = None
    if rank == 0:
        import ctypes as C
        C_double = C.c_double()
        nat.check(nat.lib().usvgpu_fp64_peak(local_rank, 148 * 8, 20000, C.byref(C_double), None))
        dfma = float(C_double.value)

    # ---- driver-visible evidence for the other BASELINE.json configs (CUDA-event timed, not the headline) ---------
    extra = {}
    if not args.no_extra and args.env == "usv-asmc-simple" and not args.no_obstacles:
        def side_run(tag, env_id, n_envs, no_obst, f32, batches, steps, label):
            slo, shi = shard_range(n_envs * world if tag != "configs3" else n_envs, rank, world)
            w2 = Workload(usv, torch, dev, local_rank, shi - slo, slo, args, env_id, no_obst, f32, batches, args.seed + 100)
            w2.preroll(8)
            g2 = w2.capture(steps)
            g2.replay()
            msx, _, _ = timed_blocks(torch, dist, dev, world, g2.replay, 3)
            tt = torch.tensor([float(shi - slo)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt)
            total = float(tt.item())
            m = statistics.median(msx)
            w2.close()
            extra[tag] = {"workload": label, "envs_total": int(total), "steps": steps, "ms_per_step": m / steps,
                          "value": total * steps / (m * 1e-3), "unit": UNIT}

        side_run("configs1", "usv-asmc-simple", 4096, True, args.obstacles_f32, 1, 200,
                 "BASELINE.json configs[1]: 4096 lock-step envs per GPU, path following without obstacles (ignore_obstacles), ASMC")
        side_run("configs1_legacy", "usv-asmc-v0", 4096, True, args.obstacles_f32, 1, 200,
                 "configs[1] secondary reading: usv-asmc-v0 (UsvAsmcEnv, one sub-step per step), 4096 envs per GPU")
        side_run("configs3", "usv-asmc-simple", 1048576, False, args.obstacles_f32, 1, 12,
                 f"BASELINE.json configs[3]: 1 048 576 envs with obstacles in total, sharded over {world} GPU(s) by env index range")
        side_run("f64_obstacle_table", "usv-asmc-simple", args.envs_per_gpu, False, 0, args.batches, K,
                 "the headline workload with the API's default float64 obstacle table (896-byte sensor-input rows)")

    clocks = sampler.stop()
    if rank == 0:
        peak, peak_src = measured_peaks()
        abytes = algorithmic_bytes(not args.no_obstacles, bool(args.obstacles_f32))
        survey_bytes = SURVEY_BYTES_F64_STATE["obstacles" if not args.no_obstacles else "no_obstacles"]
        step_s = (ms_block * 1e-3) / K
        kb = kernel_bytes(not args.no_obstacles, bool(args.obstacles_f32), args.env == "usv-simple")
        if args.env == "usv-simple":
            abytes -= 2 * 14 * 8
        if legacy:
            abytes = survey_bytes = legacy_bytes(args.env)
            kb = {legacy_kernel(args.env): abytes, "-": 0}
        elif args.env == "usv-asmc-simple":
            # SURVEY.md 8d's bytes shared out over the two kernels: the sensor kernel moves the obstacle table, n_obs,
            # the observation row, reward and flags (966 B with obstacles, 578 B without), the dynamics kernel the rest;
            # the hand-over between them exists only because the step is two kernels and is not counted
            sens = (32 * 3 * (4 if args.obstacles_f32 else 8) + 4 if not args.no_obstacles else 0) + 143 * 4 + 4 + 2
            kb = {"usv::dynamics_kernel": survey_bytes - (966 if not args.no_obstacles else 578), "usv::sensor_kernel": sens}
        knames = list(kb)
        dom = 0 if kms[0] >= kms[1] else 1
        dom_s = kms[dom] * 1e-3
        achieved = kb[knames[dom]] * n / dom_s / 1e9
        nm = ncu_metrics()
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": ms_block / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": dtype_of(args),
            "data": "synthetic", "config": workload_config(args, world),
            "timing": {"mode": "the K-step block replayed as one CUDA graph (one host call per block; every step is its own "
                               "kernel launches inside it); CUDA events around the block, barrier + synchronize on both sides, max "
                               "over ranks, median of `repeats` blocks",
                       "repeats": R, "block_ms_max_over_ranks": ms_graph, "block_ms_rank0": ms_graph_own,
                       "host_enqueue_us_per_step": 1e6 * statistics.median(host_graph) / K,
                       "python_loop": {"value": value_loop, "block_ms_max_over_ranks": ms_loop,
                                       "host_enqueue_us_per_step": 1e6 * statistics.median(host_loop) / K,
                                       "what": "the same K steps as K VecEnv.step_tensor() calls (Python -> ctypes -> launch)"}},
            "sub_steps_per_sec": value * (1 if (legacy or args.env == "usv-simple") else 20),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "variants": e2e_variants,
                    "api": "UsvGpuVecEnv.step(np.ndarray) -> np.ndarray obs/rew/dones + infos of done envs "
                           "(usvgpu_step_host: page-locked host buffers, H2D + kernels + D2H + sync per step)"},
            "gpu_launches": launches_per_step * K,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(knames[dom], n), "peak_source": peak_src, "kernel": knames[dom],
                         "algorithmic_bytes_per_env_step": kb[knames[dom]],
                         "bytes_source": ("SURVEY.md 8d (float64 state), shared out over the kernels without the hand-over; "
                                          "kernels_moved_bytes_per_env_step counts what each kernel really moves")
                         if args.env == "usv-asmc-simple" else "DESIGN.md 6",
                         "kernels_moved_bytes_per_env_step": kernel_bytes(not args.no_obstacles, bool(args.obstacles_f32),
                                                                          args.env == "usv-simple") if not legacy else None,
                         "launch_us": dom_s * 1e6,
                         "kernels_us": {knames[0]: kms[0] * 1e3, knames[1]: kms[1] * 1e3},
                         "kernels_bytes_per_env_step": kb,
                         "whole_step": {"algorithmic_bytes_per_env_step": survey_bytes, "own_count_bytes": abytes,
                                        "us": step_s * 1e6, "achieved": survey_bytes * n / step_s / 1e9,
                                        "frac": survey_bytes * n / step_s / 1e9 / peak},
                         "binding_resource": "issue slots + FP64 pipe, not HBM: see roofline_fp64 (float64 is the "
                                             "reference's precision and parity needs it, DESIGN.md 5)"},
            "targets": {"north_star_env_steps_per_s_8gpu": 1e10, "north_star_hbm_frac": 0.60,
                        "per_gpu_needed": 1.25e9, "per_gpu_measured": value / world,
                        "met_throughput_per_gpu": bool(value / world >= 1.25e9),
                        "met_hbm_frac": bool(survey_bytes * n / step_s / 1e9 / peak >= 0.60),
                        "note": "per-GPU throughput grows with the shard: 1.27e9 at 1 048 576 envs per GPU, 1.012e10 on "
                                "8 GPUs x 1 048 576 envs (profiles/r02ac_bench_8gpu_1M_per_gpu.json; --envs-per-gpu 1048576)"},
            "clocks": clocks,
            "episode_stats": stats,
            "episode_mix": {k: (stats[k] / max(stats["episodes"], 1)) for k in ("terminated", "out_of_bounds", "time_limit")},
        }
        if dfma:
            # FP64 roofline: warp-level FP64 instructions per env-step (ncu, profiles/ncu_metrics.json) x envs / step time
            # against the measured DFMA issue rate of this GPU (usvgpu_fp64_peak, timed in this run)
            wi = nm.get("fp64_warp_inst_per_env_step")
            rf = {"bound": "fp64_pipe", "peak_dfma_per_s": dfma, "peak_warp_inst_per_s": dfma / 32,
                  "fp64_warp_inst_per_env_step": wi, "source": nm.get("source")}
            if wi:
                rf["achieved_warp_inst_per_s"] = wi * n / step_s
                rf["frac"] = wi * n / step_s / (dfma / 32)
            line["roofline_fp64"] = rf
            line["issue_slot_util"] = nm.get("issue_slot_util")
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu_baseline:
            threads = host_threads()
            v, st, el = cpu_oracle_rate(args.ref_envs, args.cpu_budget_s, threads, bool(args.no_obstacles), args.seed,
                                        legacy, args.env)
            v1, st1, el1 = cpu_oracle_rate(min(args.ref_envs, 1024), max(2.0, args.cpu_budget_s / 4), 1,
                                           bool(args.no_obstacles), args.seed, legacy, args.env)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{args.ref_envs} lock-step envs x {st} steps ({el:.1f} s), C oracle, OpenMP",
                                    "single_core_value": v1}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _json_only_stdout():
    """The contract is ONE JSON line on stdout: native libraries that write to file descriptor 1 (NCCL prints its
    version banner there) are sent to stderr, Python's sys.stdout keeps the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--repeats", type=int, default=5, help="timed K-step blocks; the median block is reported")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--env", default="usv-asmc-simple", choices=["usv-asmc-simple", "usv-simple", "usv-asmc-v0", "usv-asmc-ye-int-v0", "usv-pid-v0"])
    ap.add_argument("--envs-per-gpu", type=int, default=65536)
    ap.add_argument("--batches", type=int, default=3)
    ap.add_argument("--preroll", type=int, default=480, help="untimed steps per batch before warm-up (episode mixing)")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[1] / configs[3] / f64-table side measurements")
    ap.add_argument("--e2e-all", action="store_true", help="also time the mapped-observation variant of the host face")
    ap.add_argument("--no-obstacles", action="store_true", help="configs[1] semantics (ignore_obstacles=True)")
    ap.add_argument("--obstacles-f32", type=int, default=1)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=60)
    ap.add_argument("--mapped-obs", type=int, default=0, help="e2e: sensor kernel stores obs rows straight to mapped host memory")
    ap.add_argument("--ref-envs", type=int, default=4096, help="bounded CPU sample: lock-step envs per oracle step")
    ap.add_argument("--cpu-budget-s", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        _json_only_stdout()
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: spawn one rank per GPU ourselves
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29400 + os.getpid() % 500), os.path.abspath(__file__)]
        raise SystemExit(subprocess.call(cmd + sys.argv[1:]))
    _json_only_stdout()
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
