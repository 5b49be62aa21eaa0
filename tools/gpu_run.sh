NDTB_LIB=$PWD/slam_sam_b200/libndtb_old.so timeout 120 python tools/trace_hash.py 2>&1 | tail -1
timeout 120 python tools/trace_hash.py 2>&1 | tail -1
for v in _old ""; do echo "variant $v"; NDTB_LIB=$PWD/slam_sam_b200/libndtb$v.so NDTB_EVENTS=0 timeout 120 python tools/prof_step.py 5 2>&1 | tail -2 | cut -c1-140; done
