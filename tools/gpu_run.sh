mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest.log 2>&1; tail -3 gpurun_out/r1i_pytest.log
python tools/prof_step.py 3 2>&1 | tail -2
NDTB_DEBUG_TIMING=1 python tools/eval_scaling.py > gpurun_out/r1i_scaling.log 2>&1; grep -E "^n=|dbg" gpurun_out/r1i_scaling.log | awk '/dbg/{c++; if (c%23==0) print; next} {print}'
python tools/eval_scaling.py 2>&1 | grep "^n="
