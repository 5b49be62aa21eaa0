timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1
