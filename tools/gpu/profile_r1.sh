# Round-1 profiling pass (one gpurun call; every ncu run follows a plain run of the same command).
set -x
O=gpurun_out
B1="python bench.py --batch 8192 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$B1 > $O/prof_bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_bench_r1.csv $B1 > $O/ncu_bench_launches.log 2>&1
$B1 > $O/prof_bench_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_batched -s 3 -c 1 -f -o $O/prof_batched_r1 $B1 > $O/ncu_batched_full.log 2>&1
S1="python tools/bench_single.py --pivots 4 --reps 1"
$S1 > $O/prof_single_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_single -s 2 -c 1 -f -o $O/prof_single_r1 $S1 > $O/ncu_single_full.log 2>&1
G1="python tools/bench_gj.py --batch 592 --reps 1 --cpu 0 --check 0"
$G1 > $O/prof_gj_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gauss_jordan -s 1 -c 1 -f -o $O/prof_gj_r1 $G1 > $O/ncu_gj_full.log 2>&1
ls -la $O
tail -3 $O/ncu_*.log
