timeout 300 python -m pytest tests/test_gpu_session.py -m gpu -x -q -k two_handles 2>&1 | tail -6
