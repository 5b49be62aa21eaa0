mkdir -p gpurun_out
python tools/build_only.py 6 > gpurun_out/r2d_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'voxel_stats|radix_scatter|radix_hist|seg_kernel' -s 30 -c 8 -o gpurun_out/r2d_build python tools/build_only.py 6 > gpurun_out/r2d_ncu.log 2>&1
tail -8 gpurun_out/r2d_plain.log; tail -3 gpurun_out/r2d_ncu.log
