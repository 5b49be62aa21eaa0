mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_lidar_decode.py tests/test_facade_cpp.py -m gpu -x -q 2>&1 | tail -25
