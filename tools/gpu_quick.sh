python -m pytest tests/test_render.py tests/test_replay.py tests/test_video.py -x -q 2>&1 | tail -6
