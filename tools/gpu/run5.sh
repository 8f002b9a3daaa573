set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1_c.json 2> gpurun_out/bench_r1_c.err; tail -c 2600 gpurun_out/bench_r1_c.json; tail -5 gpurun_out/bench_r1_c.err
