mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_svn.py -m gpu -x -q 2>&1 | tail -3
for v in base mb4; do
  if [ $v = base ]; then unset NDTB_LIB; else export NDTB_LIB=$PWD/gpurun_dev/libndtb_mb4.so; fi
  timeout 600 python bench.py --gpus 1 --workload svn --steps 6 --warmup 2 > gpurun_out/r2u_$v.json 2> gpurun_out/r2u_$v.err; tail -1 gpurun_out/r2u_$v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r2u_$v.json').read().strip().splitlines()[-1])
print('$v', d['value'], d['svn']['eval_us_per_iter'], d['svn']['stein_us_per_iter'], d['svn']['max_abs_translation_error_m'])
PY
done
