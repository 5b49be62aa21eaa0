mkdir -p gpurun_out
echo "== dbg"; NDTB_SORT_DEBUG=1 python tools/build_only.py 4 2>&1 | grep "sort dbg" | tail -2
