python -m pytest tests -m gpu -x -q -k "other_env_kinds" 2>&1 | tail -12
