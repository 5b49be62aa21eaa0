mkdir -p gpurun_out
echo "== dbg"; NDTB_SORT_DEBUG=1 python tools/build_only.py 4 2>&1 | grep "sort dbg" | tail -2
echo "== timing"; NDTB_BUILD_TIMING=1 python tools/build_only.py 5 2>&1 | grep "build timing" | tail -2
echo "== plain"; python tools/build_only.py 6 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_pclomp.py tests/test_gpu_svn.py tests/test_voxelgrid_filter.py tests/test_window.py tests/test_multigrid.py tests/test_gpu_large_map.py -m gpu -x -q 2>&1 | tail -30
