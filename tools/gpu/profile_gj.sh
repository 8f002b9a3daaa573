set -x
O=gpurun_out
G1="python tools/bench_gj.py --batch 592 --reps 1 --cpu 0 --check 8"
$G1 > $O/prof_gj_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gauss_jordan -s 1 -c 1 -f -o $O/prof_gj_${TAG:-r1e} $G1 > $O/ncu_gj_full.log 2>&1
cat $O/prof_gj_plain.log | cut -c1-400
