python -m pytest tests/test_video.py tests/test_replay.py -x -q 2>&1 | tail -8
python -m pytest tests -m gpu -x -q -k "collect" 2>&1 | tail -3
