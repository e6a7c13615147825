timeout 300 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -s -k pivot 2>&1 | grep -E "pivot|passed|failed|assert" | head
