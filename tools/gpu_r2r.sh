mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_svn.py tests/test_gpu_session.py tests/test_gpu_baseline_configs.py tests/test_gpu_pclomp.py -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --gpus 1 --workload svn --steps 8 --warmup 3 > gpurun_out/r2r_svn1.json 2> gpurun_out/r2r_svn1.err; tail -2 gpurun_out/r2r_svn1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2r_svn1.json').read().strip().splitlines()[-1])
print('svn value',d['value'],d['svn'])
PY
timeout 900 python bench.py --no-sub-workloads --no-cpu-baseline > gpurun_out/r2r_bench.json 2> gpurun_out/r2r_bench.err; tail -3 gpurun_out/r2r_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2r_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'serial',d['e2e_serial']['value'],d['e2e_serial']['ms_per_step'],'build_ms',d['config']['build_ms'],'align_ms',d['config']['align_ms'])
print(d['e2e'].get('build_ms'), d['e2e'].get('align_ms'))
PY
