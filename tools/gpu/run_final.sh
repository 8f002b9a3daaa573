# Round-end style verification: tests, smoke, both bench arms, config 3 and 5 tools.
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 900 gpurun_out/bench_ref.json
timeout 900 python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; tail -c 3000 gpurun_out/bench_ours.json; tail -3 gpurun_out/bench_ours.err
timeout 600 python tools/bench_gj.py --reps 3 > gpurun_out/gj.json 2>&1; tail -c 900 gpurun_out/gj.json
timeout 600 python tools/bench_single.py --pivots 200 --reps 2 > gpurun_out/single.json 2>&1; tail -c 900 gpurun_out/single.json
