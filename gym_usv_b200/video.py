"""VecVideoRecorder for the GPU VecEnv (train_test/sb3_train_vec.py:69: ``VecVideoRecorder(env, f"videos/{run.id}",
record_video_trigger=video_trigger)`` sits between ``make_vec_env`` and ``VecFrameStack``).

stable_baselines3's recorder calls ``env.render()`` after every step and encodes with moviepy / ffmpeg; neither is in
this image, and there is no window system on either box.  This class keeps the recorder's contract -- trigger on the
step counter, ``video_length`` frames per clip, one file per clip named
``{name_prefix}-step-{start}-to-step-{end}`` -- with frames drawn by ``usvgpu_render`` (device buffer, tiled like SB3's
``tile_images``) and written as an animated PNG (zlib only; plays in any browser) plus, optionally, the raw frame stack
as ``.npy``.  It forwards the tensor face of the wrapped env, so ``UsvGpuVecFrameStack`` / ``collect_rollouts`` /
``collect_transitions`` work on top of it unchanged.
"""
from __future__ import annotations

import os
import struct
import zlib
from typing import Callable, List, Optional

import numpy as np
import torch


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_apng(path: str, frames: np.ndarray, fps: int = 30, level: int = 1) -> None:
    """frames uint8 [T, H, W, 3] -> animated PNG (PNG 1.2 + APNG 1.0: IHDR acTL (fcTL IDAT|fdAT)* IEND, filter 0)."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    t, h, w, c = frames.shape
    assert c == 3 and t >= 1
    out = [b"\x89PNG\r\n\x1a\n", _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)),
           _chunk(b"acTL", struct.pack(">II", t, 0))]
    seq = 0
    for k in range(t):
        out.append(_chunk(b"fcTL", struct.pack(">IIIIIHHBB", seq, w, h, 0, 0, 1, max(int(fps), 1), 0, 0)))
        seq += 1
        raw = np.concatenate([np.zeros((h, 1), np.uint8), frames[k].reshape(h, w * 3)], axis=1).tobytes()  # filter byte 0 per row
        z = zlib.compress(raw, level)
        if k == 0:
            out.append(_chunk(b"IDAT", z))
        else:
            out.append(_chunk(b"fdAT", struct.pack(">I", seq) + z))
            seq += 1
    out.append(_chunk(b"IEND", b""))
    with open(path, "wb") as f:
        f.write(b"".join(out))


class UsvGpuVecVideoRecorder:
    """stable_baselines3 ``VecVideoRecorder`` semantics over ``UsvGpuVecEnv`` (render_mode "rgb_array")."""

    def __init__(self, venv, video_folder: str, record_video_trigger: Callable[[int], bool], video_length: int = 200,
                 name_prefix: str = "rl-video", save_npy: bool = False):
        if getattr(venv, "legacy", False):
            raise NotImplementedError("the path-following envs have no SimpleEnvVisualizer to record")
        self.venv = venv
        self.video_folder = os.path.abspath(video_folder)
        os.makedirs(self.video_folder, exist_ok=True)
        self.record_video_trigger, self.video_length = record_video_trigger, int(video_length)
        self.name_prefix, self.save_npy = name_prefix, bool(save_npy)
        self.step_id = 0
        self.recording = False
        self._frames: List[torch.Tensor] = []
        self._start = 0
        self.video_paths: List[str] = []

    def __getattr__(self, name):  # everything else is the wrapped env's (tensor face, spaces, state access ...)
        return getattr(self.venv, name)

    # ---- recorder state machine (VecVideoRecorder.reset / step_wait / start_video_recorder / close_video_recorder)
    def _tiled(self) -> torch.Tensor:
        imgs = self.venv.render_tensor()
        k, hh, ww, _ = imgs.shape
        rows = int(np.ceil(np.sqrt(k)))
        cols = int(np.ceil(k / rows))
        if rows * cols != k:
            pad = torch.zeros((rows * cols - k, hh, ww, 3), dtype=torch.uint8, device=imgs.device)
            imgs = torch.cat([imgs, pad])
        return imgs.reshape(rows, cols, hh, ww, 3).permute(0, 2, 1, 3, 4).reshape(rows * hh, cols * ww, 3)

    def _start_recording(self):
        self._finish()
        self._start, self._frames, self.recording = self.step_id, [self._tiled()], True

    def _finish(self):
        if self.recording and self._frames:
            frames = torch.stack(self._frames).cpu().numpy()
            base = os.path.join(self.video_folder,
                                f"{self.name_prefix}-step-{self._start}-to-step-{self._start + self.video_length}")
            write_apng(base + ".png", frames, fps=self.venv.metadata.get("render_fps", 30))
            if self.save_npy:
                np.save(base + ".npy", frames)
            self.video_paths.append(base + ".png")
        self.recording, self._frames = False, []

    def _after_step(self):
        self.step_id += 1
        if self.recording:
            self._frames.append(self._tiled())
            if len(self._frames) > self.video_length:
                self._finish()
        elif self.record_video_trigger(self.step_id):
            self._start_recording()

    # ---- the wrapped faces
    def reset_tensor(self, *a, **k):
        obs = self.venv.reset_tensor(*a, **k)
        if self.record_video_trigger(self.step_id):
            self._start_recording()
        return obs

    def step_tensor(self, actions):
        out = self.venv.step_tensor(actions)
        self._after_step()
        return out

    def reset(self):
        obs = self.venv.reset()
        if self.record_video_trigger(self.step_id):
            self._start_recording()
        return obs

    def step_async(self, actions):
        self.venv.step_async(actions)

    def step_wait(self):
        out = self.venv.step_wait()
        self._after_step()
        return out

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self._finish()
        self.venv.close()
