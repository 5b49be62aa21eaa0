mkdir -p gpurun_out
python tools/sanitize_small.py > gpurun_out/r2z_plain.log 2>&1 || { tail -20 gpurun_out/r2z_plain.log; exit 1; }
tail -3 gpurun_out/r2z_plain.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 3 python tools/sanitize_small.py > gpurun_out/r2z_memcheck.log 2>&1; echo "memcheck rc=$?"
grep -c "Invalid\|ERROR SUMMARY" gpurun_out/r2z_memcheck.log; tail -5 gpurun_out/r2z_memcheck.log | cut -c1-200
