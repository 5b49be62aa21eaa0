# device Newton step of the resident kernel: the session tests, then the bench probe with the step on and off
mkdir -p gpurun_out
grep -m1 -o "model name.*" /proc/cpuinfo; grep -m1 -c " fma " /proc/cpuinfo
timeout 600 python -m pytest tests/test_gpu_session.py -x -q -m gpu 2>&1 | tail -15
for dn in 0 1; do
  NDTB_BENCH_DEVICE_NEWTON=$dn timeout 600 python bench.py --no-sub-workloads --no-cpu-baseline > gpurun_out/newton_$dn.json 2> gpurun_out/newton_$dn.err; tail -2 gpurun_out/newton_$dn.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/newton_$dn.json').read().strip().splitlines()[-1])
print('device_newton=$dn value',d['value'],'build',d['config']['build_ms'],'align',d['config']['align_ms'],'sess_eval',d['roofline']['session_eval_us'],'evals',d['config']['evaluations_per_align'])
print('  e2e',d['e2e']['value'],'align',d['e2e']['align_ms'],'stats',{k:v for k,v in d['device_newton'].items() if k!='what'}, 'parity', d.get('parity'))
PY
done
