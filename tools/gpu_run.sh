timeout 600 python -m pytest tests/test_voxelgrid_filter.py tests/test_gpu_pclomp.py -x -q -m gpu 2>&1 | tail -15
