python tools/stride_probe.py 2>&1 | tail -7
timeout 900 python -m pytest tests/test_gpu_pclomp.py tests/test_window.py tests/test_multigrid.py tests/test_facade_cpp.py tests/test_voxelgrid_filter.py -m gpu -x -q 2>&1 | tail -4
