timeout 900 python -m pytest tests/test_facade_cpp.py tests/test_gpu_session.py -m gpu -x -q 2>&1 | tail -6
