for v in 296 444 592 1024 1184; do
  export NDTB_MG_BLOCKS=$v
  python bench.py --workload multigrid --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('blocks=$v', d['value'], d['ms_per_step'], d['roofline']['kernel_us'], d['config']['max_abs_translation_error_m'])"
done
