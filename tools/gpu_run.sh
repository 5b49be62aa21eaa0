mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1e_pytest.log 2>&1; tail -5 gpurun_out/r1e_pytest.log
python tools/prof_step.py 3 > gpurun_out/r1e_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -s 12 -c 1 -o gpurun_out/r1e_eval python tools/prof_step.py 2 > gpurun_out/r1e_ncu2.log 2>&1
cat gpurun_out/r1e_plain.log
