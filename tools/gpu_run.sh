NDTB_NO_SESSION=1 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b.log 2>&1 && \
NDTB_NO_SESSION=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:eval_kernel_pclomp -s 60 -c 3 -o gpurun_out/r01_eval_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log | cut -c1-200
