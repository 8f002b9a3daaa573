# ncu full capture of the batched kernel on the pivot-dominated D2 distribution (one launch)
set -x
O=gpurun_out
B1="python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
$B1 > $O/prof_batched_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:simplex_batched -s 3 -c 1 -f -o $O/prof_batched_${TAG:-r1c} $B1 > $O/ncu_batched_full.log 2>&1
tail -c 600 $O/prof_batched_plain.log
