timeout 500 python -m pytest tests/test_multigpu.py -q -x -k "without_sp or native_matches" 2>&1 | tail -6
