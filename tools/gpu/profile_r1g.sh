# r1g profiling: launch list of the default bench, then ONE full capture of the batched kernel.
# Each ncu run only after the same command exited 0 without ncu.
set -x
O=gpurun_out
B1="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$B1 > $O/prof_bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_bench_r1g.csv $B1 > $O/ncu_bench_launches.log 2>&1
$B1 > $O/prof_bench_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_batched -s 3 -c 1 -f -o $O/prof_batched_r1g $B1 > $O/ncu_batched_full.log 2>&1
tail -3 $O/ncu_batched_full.log
