set -x
timeout 1500 python -m pytest tests/test_gpu_simplex.py -m gpu -x -q 2>&1 | tail -8
