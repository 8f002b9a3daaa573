# Round-1 final profiling pass: launch list of the default bench + full captures of the three kernels.
set -x
O=gpurun_out
B1="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$B1 > $O/prof_bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_bench_r1f.csv $B1 > $O/ncu_bench_launches.log 2>&1
$B1 > $O/prof_bench_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_batched -s 3 -c 1 -f -o $O/prof_batched_r1f $B1 > $O/ncu_batched_full.log 2>&1
S1="python tools/bench_single.py --pivots 4 --reps 1"
$S1 > $O/prof_single_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_single -s 2 -c 1 -f -o $O/prof_single_r1f $S1 > $O/ncu_single_full.log 2>&1
tail -c 300 $O/prof_single_plain.log
