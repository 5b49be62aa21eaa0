timeout 600 python -m pytest tests/test_gpu_session.py -m gpu -x -q -k newton 2>&1 | tail -15
