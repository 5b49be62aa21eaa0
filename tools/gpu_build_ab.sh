python -m pytest tests/test_gpu_pclomp.py tests/test_gpu_large_map.py -m gpu -x -q 2>&1 | tail -3
for v in "" 1; do
  if [ -n "$v" ]; then export NDTB_SUM_PER_VOXEL=1; else unset NDTB_SUM_PER_VOXEL; fi
  python bench.py --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('per_voxel=$v', d['value'], d['ms_per_step'], d['config']['build_ms'], d['config']['align_ms'], d['e2e']['value'], d['e2e_window']['value'], d['e2e_packets']['value'])"
done
