mkdir -p gpurun_out
echo "== plain"; python tools/build_only.py 6 2>&1 | tail -3
echo "== timing"; NDTB_BUILD_TIMING=1 python tools/build_only.py 5 2>&1 | grep "build timing" | tail -2
python tools/build_only.py 4 > gpurun_out/r2i_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'voxel_sums|voxel_finalize|bbox_kernel|radix_hist' -s 8 -c 5 -f -o gpurun_out/r2i_build python tools/build_only.py 4 > gpurun_out/r2i_ncu.log 2>&1
tail -3 gpurun_out/r2i_ncu.log
