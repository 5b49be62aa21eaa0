mkdir -p gpurun_out
python tools/prof_step.py 1 > gpurun_out/pb_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'voxel_sum|voxel_finalize|radix_scatter|gather_points|radix_hist|bbox|seg_' -c 17 -o gpurun_out/pb_build python tools/prof_step.py 1 > gpurun_out/pb_ncu.log 2>&1
tail -3 gpurun_out/pb_ncu.log
