"""VecVideoRecorder stand-in (train_test/sb3_train_vec.py:69): animated-PNG writer and the recorder's trigger logic."""
import struct
import zlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")


def read_apng(path):
    """Minimal APNG reader for the files write_apng produces: -> (frames uint8 [T, H, W, 3], fps)."""
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, w = 8, None
    frames, fps, n_frames, seqs = [], None, None, []
    while pos < len(b):
        (ln,), tag = struct.unpack(">I", b[pos:pos + 4]), b[pos + 4:pos + 8]
        data = b[pos + 8:pos + 8 + ln]
        (crc,) = struct.unpack(">I", b[pos + 8 + ln:pos + 12 + ln])
        assert crc == zlib.crc32(tag + data) & 0xFFFFFFFF, tag
        pos += 12 + ln
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", data[:10])
            assert (depth, ctype) == (8, 2)
        elif tag == b"acTL":
            n_frames, _ = struct.unpack(">II", data)
        elif tag == b"fcTL":
            seq, fw, fh, x0, y0, num, den, _, _ = struct.unpack(">IIIIIHHBB", data)
            assert (fw, fh, x0, y0) == (w, h, 0, 0)
            seqs.append(seq)
            fps = den / num
        elif tag in (b"IDAT", b"fdAT"):
            if tag == b"fdAT":
                seqs.append(struct.unpack(">I", data[:4])[0])
                data = data[4:]
            raw = np.frombuffer(zlib.decompress(data), np.uint8).reshape(h, 1 + 3 * w)
            assert not raw[:, 0].any()  # filter type 0 on every row
            frames.append(raw[:, 1:].reshape(h, w, 3))
    assert seqs == list(range(len(seqs))) and len(frames) == n_frames
    return np.stack(frames), fps


def test_apng_roundtrip(tmp_path):
    from gym_usv_b200.video import write_apng
    rng = np.random.default_rng(0)
    frames = rng.integers(0, 256, (5, 37, 53, 3), dtype=np.uint8)
    path = str(tmp_path / "clip.png")
    write_apng(path, frames, fps=30)
    got, fps = read_apng(path)
    assert np.array_equal(got, frames) and fps == 30


@pytest.mark.gpu
def test_video_recorder_trigger_and_frames(tmp_path):
    import gym_usv_b200 as usv
    from _helpers import action_stream
    venv = usv.UsvGpuVecEnv(6, device=0, seed=2, env_id="usv-simple", render_mode="rgb_array")
    venv.render_envs, venv.render_size = 6, 64
    rec = usv.UsvGpuVecVideoRecorder(venv, str(tmp_path), record_video_trigger=lambda s: s == 3, video_length=4,
                                     save_npy=True)
    env = usv.UsvGpuVecFrameStack(rec, n_stack=5)   # the order of sb3_train_vec.py:67-70
    obs = env.reset_tensor()
    assert obs.shape == (6, 5 * 143) and not rec.recording
    want = []
    for t in range(12):
        obs, rew, term, trunc = env.step_tensor(torch.tensor(action_stream(6, seed=t), device=venv.device))
        if 3 <= t + 1 <= 7:      # the frame at the trigger step, then video_length more
            want.append(venv.render())
    assert rec.step_id == 12 and not rec.recording and len(rec.video_paths) == 1
    assert rec.video_paths[0].endswith("rl-video-step-3-to-step-7.png")
    got, fps = read_apng(rec.video_paths[0])
    assert fps == 30 and got.shape == (5, 3 * 64, 2 * 64, 3)
    assert np.array_equal(got, np.stack(want))
    assert np.array_equal(np.load(rec.video_paths[0][:-4] + ".npy"), got)
    # host face + a clip cut short by close()
    rec2 = usv.UsvGpuVecVideoRecorder(venv, str(tmp_path), record_video_trigger=lambda s: s == 0, video_length=50,
                                      name_prefix="host")
    rec2.reset()
    for t in range(3):
        rec2.step(action_stream(6, seed=50 + t))
    rec2.close()
    got, _ = read_apng(rec2.video_paths[0])
    assert got.shape[0] == 4
