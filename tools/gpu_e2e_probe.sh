mkdir -p gpurun_out
timeout 600 python bench.py --no-sub-workloads --no-cpu-baseline > gpurun_out/e2e_probe.json 2> gpurun_out/e2e_probe.err; tail -2 gpurun_out/e2e_probe.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/e2e_probe.json').read().strip().splitlines()[-1])
print('resident: build',d['config']['build_ms'],'align',d['config']['align_ms'],'sess_eval',d['roofline']['session_eval_us'],'evals',d['config']['evaluations_per_align'])
print('pipelined:',d['e2e'])
print('serial:',d['e2e_serial']['ms_per_step'])
PY
