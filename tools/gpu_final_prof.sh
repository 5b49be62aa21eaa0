mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/fp_plain.json 2> gpurun_out/fp_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/fp_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/fp_ncu1.log 2>&1
NDTB_NO_SESSION=1 ncu --set full --clock-control none --import-source on -k regex:eval_kernel_pclomp -s 60 -c 3 -o gpurun_out/fp_eval python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/fp_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lidar_decode_kernel -s 3 -c 2 -o gpurun_out/fp_lidar python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/fp_ncu3.log 2>&1
ls -la gpurun_out | tail -12
