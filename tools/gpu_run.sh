NDTB_EVENTS=1 timeout 120 python tools/prof_step.py 2 > gpurun_out/plain.log 2>&1 && \
NDTB_EVENTS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:eval_kernel_pclomp -s 14 -c 1 -o gpurun_out/eval_r01d python tools/prof_step.py 2 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
