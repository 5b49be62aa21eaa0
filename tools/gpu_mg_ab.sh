for v in "" mg3 mg4; do
  if [ -n "$v" ]; then export NDTB_LIB=$PWD/gpurun_dev/libndtb_$v.so; else unset NDTB_LIB; fi
  python bench.py --workload multigrid --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', d['value'], d['ms_per_step'], d['roofline']['kernel_us'], d['config']['max_abs_translation_error_m'])"
done
