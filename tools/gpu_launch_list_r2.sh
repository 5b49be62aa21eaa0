mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_plain.json 2> gpurun_out/r2s_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r2s_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_ncu.log 2>&1
tail -2 gpurun_out/r2s_ncu.log | cut -c1-300; wc -l gpurun_out/r2s_launches.csv
