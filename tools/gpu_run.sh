timeout 600 python -m pytest tests/test_gpu_pclomp.py tests/test_gpu_session.py -x -q 2>&1 | tail -15
