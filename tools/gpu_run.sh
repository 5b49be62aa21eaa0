timeout 1200 python -m pytest tests/test_gpu_large_map.py -x -q 2>&1 | tail -25
