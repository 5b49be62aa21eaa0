mkdir -p gpurun_out
python tools/build_only.py 4 > gpurun_out/r2v_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'sort_segments|voxel_stats' -s 4 -c 4 -f -o gpurun_out/r2v_build python tools/build_only.py 4 > gpurun_out/r2v_ncu.log 2>&1
tail -2 gpurun_out/r2v_ncu.log | cut -c1-200
