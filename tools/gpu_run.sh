mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1r_pytest.log 2>&1; tail -3 gpurun_out/r1r_pytest.log
python tools/eval_scaling.py 2>&1 | grep "^n="
python tools/prof_eval_big.py 8 > gpurun_out/r1r_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:eval_kernel_pclomp -s 3 -c 1 -o gpurun_out/r1r_eval python tools/prof_eval_big.py 8 > gpurun_out/r1r_ncu.log 2>&1
