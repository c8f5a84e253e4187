"""rgb_array frames: SimpleEnvVisualizer.render_frame (gym_usv/envs/simple_env_visualizer.py:17-114).

PARITY UNPINNED against the reference's pixels: it draws with pygame, which is absent from this image, so there is no
golden frame.  What the CPU tests pin is the draw list of the reference (positions, colours, radii, widths, order,
(H, W, 3) layout) on a hand-checkable scene and the stated coverage rules against a second, independent numpy
statement; the GPU tests then require usvgpu_render to equal the C restatement bit for bit.
"""
import numpy as np
import pytest

import usv_oracle as uo

WHITE, BLUE, GREEN, RED, FRONT, OBST, PATH = ((255, 255, 255), (0, 0, 255), (0, 255, 0), (255, 0, 0), (100, 100, 0),
                                              (0, 100, 0), (100, 0, 0))


def px(frame, x, y):
    return tuple(int(c) for c in frame[y, x])  # row = canvas y (the np.transpose of simple_env_visualizer.py:112-114)


def numpy_render(position, target, dist, ox, oy, orad, ps, pe, env_hi=20.0, size=512):
    """Second statement of the coverage rules: per-pixel closed forms on index grids (no walks, no painter's loop over
    pixels) -- the rules of gym_usv_b200/csrc/usv_render.cuh written with numpy."""
    tr = lambda v: int(np.clip(v, -(1 << 20), 1 << 20))  # noqa: E731  int() truncates toward zero
    scale = size / env_hi
    yy, xx = np.mgrid[0:size, 0:size].astype(np.int64)
    frame = np.full((size, size, 3), 255, np.uint8)

    def circle(cx, cy, r, rgb):
        if r >= 1:
            frame[(2 * xx + 1 - 2 * cx) ** 2 + (2 * yy + 1 - 2 * cy) ** 2 <= 4 * r * r] = rgb

    def line(x0, y0, x1, y1, w, rgb):
        dx, dy = x1 - x0, y1 - y0
        lo, hi = -((w - 1) // 2), w // 2
        if abs(dx) >= abs(dy):
            k = (xx - x0) * (1 if dx >= 0 else -1)
            m = (2 * abs(dy) * k + abs(dx)) // (2 * abs(dx)) if dx else np.zeros_like(k)
            off = yy - (y0 + m * (1 if dy >= 0 else -1))
            ok = (k >= 0) & (k <= abs(dx))
        else:
            k = (yy - y0) * (1 if dy >= 0 else -1)
            m = (2 * abs(dx) * k + abs(dy)) // (2 * abs(dy))
            off = xx - (x0 + m * (1 if dx >= 0 else -1))
            ok = (k >= 0) & (k <= abs(dy))
        frame[ok & (off >= lo) & (off <= hi)] = rgb

    x, y, psi = position
    circle(tr(target[0] * scale), tr(target[1] * scale), 10, BLUE)
    pitch = ((2.0 / 3) * (2 * np.pi)) / len(dist)
    for k, d in enumerate(dist):
        a = (-np.pi * 2 / 3 + k * pitch) + psi
        line(tr(x * scale), tr(y * scale), tr((d * np.cos(a) + x) * scale), tr((d * np.sin(a) + y) * scale), 1, GREEN)
    circle(tr(x * scale), tr(y * scale), 10, RED)
    circle(tr((np.cos(psi) * 0.1 + x) * scale), tr((np.sin(psi) * 0.1 + y) * scale), 8, FRONT)
    for j in range(len(ox)):
        circle(tr(ox[j] * scale), tr(oy[j] * scale), tr(orad[j] * scale), OBST)
    line(tr(ps[0] * scale), tr(ps[1] * scale), tr(pe[0] * scale), tr(pe[1] * scale), 5, PATH)
    return frame


def test_oracle_frame_draw_list_on_a_hand_checked_scene():
    """512 x 512, scale 25.6 px/m (simple_env_visualizer.py:37): boat at (10, 10) heading +x, target (5, 5), one
    obstacle, 1 m rays, path along y = 2 m."""
    f = uo.render_frame([10.0, 10.0, 0.0], [5.0, 5.0], np.full(128, 1.0), [15.0], [15.0], [0.5], [2.0, 2.0], [18.0, 2.0])
    assert f.shape == (512, 512, 3) and f.dtype == np.uint8
    assert px(f, 0, 0) == WHITE and px(f, 511, 511) == WHITE                       # :35
    assert px(f, 128, 128) == BLUE and px(f, 137, 128) == BLUE and px(f, 138, 128) == WHITE  # :40-45 radius 10 at (128, 128)
    assert px(f, 250, 250) == RED                                                  # :64-69 agent at (256, 256)
    assert px(f, 258, 256) == FRONT and px(f, 265, 256) == FRONT                   # :72-79 front marker at (258, 256), radius 8
    assert px(f, 266, 256) == GREEN and px(f, 281, 256) == GREEN                   # :48-61 ray 64 points along +x, 25 px long
    assert px(f, 282, 256) == WHITE
    assert px(f, 384, 384) == OBST and px(f, 384 + 11, 384) == OBST and px(f, 384 + 12, 384) == WHITE  # :82-91 r = int(12.8)
    for y, want in ((48, WHITE), (49, PATH), (51, PATH), (53, PATH), (54, WHITE)):  # :94-100 width 5 around y = 51
        assert px(f, 300, y) == want, y
    assert px(f, 50, 51) == WHITE and px(f, 51, 51) == PATH and px(f, 460, 51) == PATH and px(f, 461, 51) == WHITE
    # painter's order: the path line is drawn last, the agent over the rays, the obstacles over the agent
    g = uo.render_frame([10.0, 10.0, 0.0], [10.0, 10.0], np.full(128, 1.0), [10.0], [10.2], [0.2], [2.0, 10.0], [18.0, 10.0])
    assert px(g, 256, 256) == PATH and px(g, 249, 250) == RED and px(g, 256, 262) == OBST and px(g, 256, 250) == FRONT
    # disc area: pixel centres inside radius 10 -> 316 pixels (Gauss circle count on the half-integer lattice)
    h = uo.render_frame([-50.0, -50.0, 0.0], [10.0, 10.0], np.zeros(128), [], [], [], [-60.0, -60.0], [-70.0, -60.0])
    assert int((h == np.array(BLUE, np.uint8)).all(-1).sum()) == 316


@pytest.mark.parametrize("size", [64, 100, 512])
def test_oracle_frame_equals_numpy_statement_on_random_scenes(size):
    rng = np.random.default_rng(size)
    for trial in range(6 if size < 512 else 3):
        nob = int(rng.integers(0, 30))
        pos = np.array([rng.uniform(-1, 21), rng.uniform(-1, 21), rng.uniform(-10, 10)])
        tgt = rng.uniform(0, 20, 2)
        dist = np.where(rng.random(128) < 0.5, 100.0, rng.uniform(0, 30, 128))
        ox, oy, orad = rng.uniform(0, 20, nob), rng.uniform(0, 20, nob), rng.uniform(0.01, 0.5, nob)
        ps = rng.normal(10, 0.5, 2)
        ang = rng.uniform(-np.pi, np.pi)
        pe = ps + 105 * np.array([np.cos(ang), np.sin(ang)])
        a = uo.render_frame(pos, tgt, dist, ox, oy, orad, ps, pe, size=size)
        b = numpy_render(pos, tgt, dist, ox, oy, orad, ps, pe, size=size)
        assert np.array_equal(a, b), (size, trial, int((a != b).any(-1).sum()))
        assert (a == np.array(GREEN, np.uint8)).all(-1).sum() > 0


def test_oracle_frame_degenerate_inputs():
    """Zero-length rays, a point-like path, NaN pose, everything off canvas: no crash, nothing out of bounds."""
    f = uo.render_frame([np.nan, 5.0, 0.0], [1e9, -1e9], np.zeros(128), [5.0], [5.0], [0.01], [3.0, 3.0], [3.0, 3.0], size=64)
    assert f.shape == (64, 64, 3)
    assert px(f, 9, 9) == PATH and px(f, 9, 7) == PATH and px(f, 9, 6) == WHITE  # the point-like path: 1 x 5 pixels


# ------------------------------------------------------------------------------------------ GPU
torch = pytest.importorskip("torch")


def _oracle_frames(env, ov, obs, idx, size):
    st = env.get_state()
    frames = []
    for i in idx:
        ps, pe = st["epi"][0:2, i], st["epi"][2:4, i]
        if st["steps"][i] == 0:  # the random target of reset() (simple_env.py:245); the GPU keeps it as float32
            tgt = ov.envs["target"][i].astype(np.float32).astype(np.float64)
        else:
            a = st["dyn"][22, i]
            tgt = np.array([ps[0] + a * (pe[0] - ps[0]), ps[1] + a * (pe[1] - ps[1])])
        nob = int(st["n_obs"][i])
        ob = st["obstacles"][i].astype(np.float64)
        dist = obs[i, 15:].astype(np.float64) * 100.0
        frames.append(uo.render_frame(st["dyn"][0:3, i], tgt, dist, ob[0, :nob], ob[1, :nob], ob[2, :nob], ps, pe, size=size))
    return np.stack(frames)


@pytest.mark.gpu
@pytest.mark.parametrize("env_id,obstacles_f32,size", [("usv-asmc-simple", True, 512), ("usv-asmc-simple", False, 100),
                                                       ("usv-simple", True, 64), ("usv-asmc-simple", True, 257)])
def test_gpu_frames_equal_the_cpu_restatement(env_id, obstacles_f32, size):
    import gym_usv_b200 as usv
    from _helpers import action_stream
    N = 96
    env = usv.UsvGpuVecEnv(N, device=0, seed=9, env_id=env_id, obstacles_f32=obstacles_f32, render_mode="rgb_array")
    ov = uo.OracleVecEnv(N, seed=9, kinematic=env_id == "usv-simple", obstacles_f32=obstacles_f32,
                         max_episode_steps=500 if env_id == "usv-simple" else 1000)
    obs = env.reset_tensor()
    ov.reset()
    idx = [0, 1, 5, 17, 95]
    got = env.render_tensor(idx, size=size).cpu().numpy()   # freshly reset envs: random target, stale (zero) rays
    want = _oracle_frames(env, ov, obs.cpu().numpy(), idx, size)
    assert np.array_equal(got, want), int((got != want).any(-1).sum())
    for t in range(60):
        a = action_stream(N, seed=100 + t)
        obs, *_ = env.step_tensor(torch.tensor(a, device=env.device))
        ov.step(a)
    idx = list(range(0, N, 7))
    got = env.render_tensor(idx, size=size).cpu().numpy()
    want = _oracle_frames(env, ov, obs.cpu().numpy(), idx, size)
    assert got.shape == (len(idx), size, size, 3)
    assert np.array_equal(got, want), int((got != want).any(-1).sum())
    assert (got == np.array(GREEN, np.uint8)).all(-1).sum() > 100 and (got == np.array(OBST, np.uint8)).all(-1).sum() > 100
    # default face: the first render_envs envs; SB3-style tiled picture
    env.render_envs, env.render_size = 6, size
    imgs = env.get_images()
    assert len(imgs) == 6 and np.array_equal(imgs[0], got[0])
    if size == 64:
        tiled = env.render()
        assert tiled.shape == (3 * 64, 2 * 64, 3) and np.array_equal(tiled[:64, :64], imgs[0])
        assert np.array_equal(tiled[64:128, 64:128], imgs[3])
    env.close()


@pytest.mark.gpu
def test_gpu_render_errors_are_loud():
    import gym_usv_b200 as usv
    env = usv.UsvGpuVecEnv(8, device=0, seed=1, env_id="usv-asmc-v0")
    with pytest.raises(NotImplementedError):
        env.render_tensor()
    env.close()
    env = usv.UsvGpuVecEnv(8, device=0, seed=1)
    env.reset_tensor()
    with pytest.raises(IndexError):
        env.render_tensor([8])
    with pytest.raises(RuntimeError):
        env.render_tensor([0], size=-1)
    # through the C ABI an index outside the shard leaves its frame untouched instead of reading out of bounds
    import ctypes as C
    from gym_usv_b200 import _native as nat
    idx = torch.tensor([0, 99, -5], dtype=torch.int64, device=env.device)
    out = torch.full((3, 64, 64, 3), 7, dtype=torch.uint8, device=env.device)
    nat.check(nat.lib().usvgpu_render(env._h, C.c_void_p(idx.data_ptr()), 3, C.c_void_p(env.obs.data_ptr()), 64,
                                      C.c_void_p(out.data_ptr()), None))
    torch.cuda.synchronize()
    assert (out[1:] == 7).all() and (out[0] != 7).any()
    env.close()
