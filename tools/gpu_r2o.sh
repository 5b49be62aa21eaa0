mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pclomp.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py --no-sub-workloads --no-cpu-baseline > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; tail -3 gpurun_out/r2o_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2o_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'serial',d['e2e_serial']['value'],d['e2e_serial']['ms_per_step'],'build_ms',d['config']['build_ms'],'align_ms',d['config']['align_ms'])
PY
