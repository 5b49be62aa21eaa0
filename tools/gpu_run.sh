timeout 600 python -m pytest tests/test_real_scan_pair.py -x -q 2>&1 | tail -12
