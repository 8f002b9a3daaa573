set -x
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -q 2>&1 | tail -15
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_sharded.py 2>&1 | tail -5
timeout 300 python tools/bench_single.py --pivots 50 --reps 1 2>&1 | tail -2
