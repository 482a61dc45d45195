timeout 600 python -m pytest tests/test_spmm_gpu.py -q -m gpu -k "tile" 2>&1 | grep -E "^E|passed|failed|PASS|FAIL" | head -40
