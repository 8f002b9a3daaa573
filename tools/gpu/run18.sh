# partial-pivoting GJ: tests + config-3 bench (reference rule must not have slowed down)
set -x
timeout 1200 python -m pytest tests/test_gpu_gauss_jordan.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/bench_gj.py --reps 3 2>&1 | tail -c 1200
