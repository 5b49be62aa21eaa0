# phase timing of the device Newton step (library built with -DNDTB_NEWTON_TIMING as slam_sam_b200/libndtb_timing.so)
mkdir -p gpurun_out
NDTB_LIB=$PWD/slam_sam_b200/libndtb_timing.so timeout 600 python bench.py --no-sub-workloads --no-cpu-baseline > gpurun_out/newton_t.json 2> gpurun_out/newton_t.err
grep "newton timing" gpurun_out/newton_t.err | tail -3
