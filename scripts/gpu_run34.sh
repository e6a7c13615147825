timeout 600 python -m pytest tests/test_gpu_helpers.py -m gpu -x -q -k "forward_backward or fused_epilogue" 2>&1 | tail -5
