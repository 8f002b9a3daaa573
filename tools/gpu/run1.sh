set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1_a.json 2> gpurun_out/bench_r1_a.err; tail -c 3000 gpurun_out/bench_r1_a.json; tail -5 gpurun_out/bench_r1_a.err
python tools/bench_single.py --pivots 100 --reps 1 > gpurun_out/single_r1_a.json 2>&1; tail -c 1500 gpurun_out/single_r1_a.json
python tools/bench_gj.py --reps 2 > gpurun_out/gj_r1_a.json 2>&1; tail -c 1500 gpurun_out/gj_r1_a.json
