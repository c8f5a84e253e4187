"""frames/s of usvgpu_render (SimpleEnvVisualizer.render_frame as a CUDA kernel) beside the CPU restatement.

python tools/bench_render.py [--frames 1024] [--size 512] [--iters 20]
One JSON line: device-resident frames/s (CUDA events), the HBM write roofline (size*size*3 bytes per frame), the
same through a D2H copy into page-locked memory, and the C oracle's frames/s on one host core.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np
    import torch
    import gym_usv_b200 as usv
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--cpu-frames", type=int, default=32)
    args = ap.parse_args()
    n = max(args.frames, 4096)
    env = usv.UsvGpuVecEnv(n, device=0, seed=0, render_mode="rgb_array")
    obs = env.reset_tensor()
    gen = torch.Generator(device=env.device).manual_seed(1)
    for _ in range(50):
        a = torch.rand((n, 2), generator=gen, device=env.device) * torch.tensor([0.8, 2.0], device=env.device) \
            + torch.tensor([0.2, -1.0], device=env.device)
        obs, *_ = env.step_tensor(a.contiguous())
    idx = torch.arange(args.frames, device=env.device)
    out = torch.empty((args.frames, args.size, args.size, 3), dtype=torch.uint8, device=env.device)
    for _ in range(3):
        env.render_tensor(idx, size=args.size, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        env.render_tensor(idx, size=args.size, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    fps = args.frames / (ms * 1e-3)
    host = torch.empty(out.shape, dtype=torch.uint8, pin_memory=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        env.render_tensor(idx, size=args.size, out=out)
        host.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
    e2e = 3 * args.frames / (time.perf_counter() - t0)
    # CPU restatement on the same scenes
    import usv_oracle as uo
    from test_render import _oracle_frames
    ov = uo.OracleVecEnv(8, seed=0)  # only consulted for envs at step 0; none here after 50 steps unless just reset
    st_steps = env.get_state(["steps"])["steps"]
    cidx = [int(i) for i in np.flatnonzero(st_steps[:args.frames] > 0)[:args.cpu_frames]]
    ob = obs.cpu().numpy()
    t0 = time.perf_counter()
    want = _oracle_frames(env, ov, ob, cidx, args.size)
    cpu_s = time.perf_counter() - t0  # includes the get_state copy: subtract a second call without frames
    t0 = time.perf_counter()
    _oracle_frames(env, ov, ob, cidx[:1], args.size)
    cpu_s = max(cpu_s - (time.perf_counter() - t0), 1e-9)
    same = bool(np.array_equal(out[cidx].cpu().numpy(), want))
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa: BLE001
        peak = 6650.0
    gbs = args.frames * args.size * args.size * 3 / (ms * 1e-3) / 1e9
    print(json.dumps({"metric": "frames_per_sec", "value": fps, "frames": args.frames, "size": args.size,
                      "ms_per_launch": ms, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s",
                                                        "frac": gbs / peak, "bytes_per_frame": args.size * args.size * 3},
                      "e2e_frames_per_sec_pinned_host": e2e,
                      "cpu_oracle_frames_per_sec_1core": (len(cidx) - 1) / cpu_s if len(cidx) > 1 else None,
                      "frames_equal_cpu_restatement": same}))


if __name__ == "__main__":
    main()
