timeout 120 python tools/trace_hash.py 2>&1 | tail -1
NDTB_EVENTS=0 timeout 120 python tools/prof_step.py 5 2>&1 | tail -3 | cut -c1-150
