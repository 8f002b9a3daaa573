# typed-A path: tests + bench with e2e_compact
set -x
timeout 900 python -m pytest tests/test_gpu_simplex.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py > gpurun_out/bench_r15.json 2> gpurun_out/bench_r15.err; tail -c 3500 gpurun_out/bench_r15.json; tail -3 gpurun_out/bench_r15.err
