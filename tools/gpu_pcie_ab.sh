# A/B of the PCIe side of an evaluation: results published by one warp as 16-byte stores + 16-byte mailbox loads (libndtb.so)
# against the 46 owner lanes storing 8-byte halves + 8-byte mailbox loads (libndtb_old_results.so, -DNDTB_RESULTS_PER_LANE)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_session.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
for rep in 1 2; do
for lib in libndtb.so libndtb_old_results.so; do
  NDTB_LIB=$PWD/slam_sam_b200/$lib timeout 600 python bench.py --no-sub-workloads --no-cpu-baseline --steps 40 > gpurun_out/pcie_$lib.json 2> gpurun_out/pcie_$lib.err; tail -1 gpurun_out/pcie_$lib.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/pcie_$lib.json').read().strip().splitlines()[-1])
ev=d['config']['evaluations_per_align']
print('$lib value',round(d['value'],1),'align_ms',round(d['config']['align_ms'],4),'sess_eval_us',round(d['roofline']['session_eval_us'],2),'overhead_us_per_eval',round(d['config']['align_ms']*1e3/ev-d['roofline']['session_eval_us'],2),'e2e',round(d['e2e']['value'],1),'kernel_us',round(d['roofline']['kernel_us'],2))
PY
done
done
