"""Device replay buffer for SAC (train_test/sb3_train_vec.py:72, config_sac) against the numpy restatement of SB3's
ReplayBuffer (oracle/sb3_replay.py; parity unpinned against SB3 itself, which is absent here)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import sb3_replay  # noqa: E402


def test_sb3_replay_restatement_wraps_and_masks_timeouts():
    rb = sb3_replay.ReplayBuffer(buffer_size=6, n_envs=2, obs_dim=3, act_dim=1)
    assert rb.buffer_size == 3
    for t in range(4):
        obs = np.full((2, 3), t, np.float32)
        new = obs + 1
        infos = [{}, {"TimeLimit.truncated": t == 1, "terminal_observation": np.full(3, -7, np.float32)}]
        sb3_replay.store_transition(rb, obs, new, np.full((2, 1), 0.5, np.float32), np.array([1.0, 2.0], np.float32),
                                    np.array([False, t in (1, 2)]), infos)
    assert rb.full and rb.pos == 1 and rb.observations[0, 0, 0] == 3  # slot 0 overwritten by step 3
    s = rb.get_samples(np.array([1, 2, 1]), np.array([1, 1, 0]))
    assert np.all(s["next_observations"][0] == -7) and np.all(s["next_observations"][2] == 2)
    assert s["dones"].ravel().tolist() == [0.0, 1.0, 0.0]  # step 1 was a time-limit truncation: masked; step 2 a real end


torch = pytest.importorskip("torch")


def _host_infos(term, trunc, tobs):
    infos = []
    for i in range(len(term)):
        d = {}
        if term[i] or trunc[i]:
            d["terminal_observation"] = tobs[i].copy()
            d["TimeLimit.truncated"] = bool(trunc[i] and not term[i])
        infos.append(d)
    return infos


@pytest.mark.gpu
@pytest.mark.parametrize("env_id,n_stack,limit", [("usv-simple", 0, 7), ("usv-asmc-simple", 0, 11), ("usv-simple", 3, 9),
                                                  ("usv-asmc-v0", 0, 5)])
def test_device_replay_buffer_equals_sb3_restatement(env_id, n_stack, limit):
    import gym_usv_b200 as usv
    N, T = 48, 40
    venv = usv.UsvGpuVecEnv(N, device=0, seed=3, env_id=env_id, max_episode_steps=limit)
    env = usv.UsvGpuVecFrameStack(venv, n_stack=n_stack) if n_stack else venv
    obs_dim = venv.obs_dim * max(n_stack, 1)
    buf = usv.DeviceReplayBuffer(buffer_size=13 * N, n_envs=N, obs_dim=obs_dim, act_dim=venv.act_dim, device=venv.device)
    ref = sb3_replay.ReplayBuffer(13 * N, N, obs_dim, venv.act_dim)
    assert buf.buffer_size == ref.buffer_size == 13
    last = env.reset_tensor().clone()
    rng = np.random.default_rng(0)
    lo, hi = venv.action_space.low, venv.action_space.high
    n_done = n_timeouts = 0
    for t in range(T):
        a = rng.uniform(lo, hi, (N, venv.act_dim)).astype(np.float32)
        at = torch.tensor(a, device=venv.device)
        new, rew, term, trunc = env.step_tensor(at)
        tobs = env.terminal_stack if n_stack else venv.terminal_obs
        buf.add(last, new, at, rew, term, trunc, tobs)
        th, uh = term.cpu().numpy().astype(bool), trunc.cpu().numpy().astype(bool)
        sb3_replay.store_transition(ref, last.cpu().numpy(), new.cpu().numpy(), a, rew.cpu().numpy(), th | uh,
                                    _host_infos(th, uh, tobs.cpu().numpy()))
        n_done += int((th | uh).sum())
        n_timeouts += int((uh & ~th).sum())
        last = new.clone()
    assert n_done > 0 and n_timeouts > 0
    assert (buf.pos, buf.full) == (ref.pos, ref.full) == (T % 13, True)
    for name in ("observations", "next_observations", "actions", "rewards"):
        assert np.array_equal(getattr(buf, name).cpu().numpy(), getattr(ref, name)), name
    assert np.array_equal(buf.dones.cpu().numpy(), ref.dones.astype(np.uint8))
    assert np.array_equal(buf.timeouts.cpu().numpy(), ref.timeouts.astype(np.uint8))
    bi, ei = rng.integers(0, 13, 300), rng.integers(0, N, 300)
    got = buf.sample(300, batch_inds=bi, env_inds=ei)
    want = ref.get_samples(bi, ei)
    for name in ("observations", "actions", "next_observations", "dones", "rewards"):
        assert np.array_equal(getattr(got, name).cpu().numpy(), want[name]), name
    bad = buf.sample(3, batch_inds=[0, 13, -1], env_inds=[N, 0, 0])  # outside the buffer: zeros, no fault
    assert not bad.observations.any() and not bad.rewards.any() and not bad.dones.any()
    g = torch.Generator(device=venv.device).manual_seed(5)
    s = buf.sample(256, generator=g)
    assert s.observations.shape == (256, obs_dim) and s.dones.shape == (256, 1) and s.actions.shape == (256, venv.act_dim)
    venv.close()


@pytest.mark.gpu
def test_collect_transitions_device_resident():
    import gym_usv_b200 as usv
    N = 256
    venv = usv.UsvGpuVecEnv(N, device=0, seed=8, env_id="usv-simple")
    env = usv.UsvGpuVecFrameStack(venv, n_stack=5)
    buf = usv.DeviceReplayBuffer(400000, N, 5 * venv.obs_dim, venv.act_dim, venv.device)  # config_sac buffer_size
    assert buf.buffer_size == 400000 // N
    obs = env.reset_tensor()
    g = torch.Generator(device=venv.device).manual_seed(0)
    obs = usv.collect_transitions(env, None, buf, obs, 8, generator=g)  # warm-up: uniform actions (learning_starts)
    actor = torch.nn.Sequential(torch.nn.Linear(5 * venv.obs_dim, 2), torch.nn.Tanh()).to(venv.device)
    obs = usv.collect_transitions(env, actor, buf, obs, 8)               # train_freq = 8
    assert buf.pos == 16 and not buf.full
    o, nx, d = buf.observations[:16], buf.next_observations[:16], buf.dones[:16].bool()
    live = ~d[:-1]
    assert torch.equal(nx[:-1][live], o[1:][live])     # the next observation of a live env is its observation one step on
    a = buf.actions[:16]
    assert float(a[..., 0].min()) >= 0.0 and float(a[..., 1].abs().max()) <= 1.0  # inside the action box of simple_env.py:27
    s = buf.sample(256, generator=g)
    assert torch.isfinite(s.observations).all() and torch.isfinite(s.rewards).all()
    venv.close()
