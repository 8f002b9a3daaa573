# Round-end style verification on one B200 (run as: gpurun --timeout 2400 -- 'bash tools/gpu/run_final.sh').
# Everything lands in gpurun_out/; copy what should be judged into profiles/.
set -x
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; tail -c 600 $O/bench_ref.json
timeout 900 python bench.py > $O/bench_ours.json 2> $O/bench_ours.err; tail -c 3000 $O/bench_ours.json; tail -3 $O/bench_ours.err
timeout 600 python tools/bench_gj.py --reps 3 > $O/gj.json 2>&1; tail -c 900 $O/gj.json
timeout 600 python tools/bench_single.py --pivots 200 --reps 2 > $O/single.json 2>&1; tail -c 900 $O/single.json
timeout 600 python tools/bench_course.py > $O/course.json 2> $O/course.err; tail -c 700 $O/course.json
timeout 300 python tools/gpu/rule_rates.py 2>&1 | tail -6
