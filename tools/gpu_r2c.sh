# round 2: new voxel-map build (device-side geometry / radix plan, 11-bit digits, look-back segments, fused gather+sums+finalize)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pclomp.py tests/test_gpu_svn.py tests/test_voxelgrid_filter.py tests/test_window.py -m gpu -x -q > gpurun_out/r2c_t1.log 2>&1; tail -12 gpurun_out/r2c_t1.log
timeout 900 python -m pytest tests/test_gpu_large_map.py tests/test_multigrid.py tests/test_gpu_baseline_configs.py tests/test_real_scan_pair.py tests/test_gpu_session.py -m gpu -x -q > gpurun_out/r2c_t2.log 2>&1; tail -12 gpurun_out/r2c_t2.log
timeout 600 python bench.py --no-cpu-baseline --no-sub-workloads > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; tail -3 gpurun_out/r2c_bench.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2c_bench.json').read().strip().splitlines()[-1])
    print('value',d['value'],'ms',d['ms_per_step'],'build_ms',d['config']['build_ms'],'align_ms',d['config']['align_ms'],'e2e',d['e2e']['value'], d['roofline_build'])
except Exception as e: print('no line',e)
PY
timeout 600 python bench.py --no-cpu-baseline --no-sub-workloads --steps 2 --warmup 3 > gpurun_out/r2c_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2c_launches.csv python bench.py --no-cpu-baseline --no-sub-workloads --steps 2 --warmup 3 > gpurun_out/r2c_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/r2c_launches.csv
