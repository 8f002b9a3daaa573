# ncu evidence for one kernel (one GPU; each ncu run only after the same command exited 0 without ncu).
#   bash tools/gpu/profile.sh batched|single|gj|verify <tag>      e.g.  ... batched r1g
# Writes gpurun_out/launches_<kernel>_<tag>.csv (per-launch gpu__time_duration) and
# gpurun_out/prof_<kernel>_<tag>.ncu-rep (--set full, source imported); summarise the report here with
#   python tools/ncu_summary.py gpurun_out/prof_<kernel>_<tag>.ncu-rep profiles/<tag>_<kernel>_ncu_summary.csv
#   python tools/ncu_facts.py <kernel key> gpurun_out/prof_<kernel>_<tag>.ncu-rep --units <pivots or systems per launch> --summary profiles/<tag>_<kernel>_ncu_summary.csv
set -x
K=${1:-batched}; TAG=${2:-r1}; O=gpurun_out
case $K in
  batched) CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-single --no-gj"; RE=simplex_batched; SKIP=3;;
  single)  CMD="python tools/bench_single.py --pivots 4 --reps 1"; RE=simplex_single; SKIP=2;;
  gj)      CMD="python tools/bench_gj.py --batch 592 --reps 1 --cpu 0 --check 8"; RE=gauss_jordan; SKIP=1;;
  verify)  CMD="python tools/gpu/verify_rate.py"; RE=verify_bounds; SKIP=1;;
  *) echo "unknown kernel $K"; exit 2;;
esac
python -c "import bench; print(bench.source_sha16())" > $O/prof_${K}_${TAG}.sha   # ties the capture to the build (tools/ncu_facts.py)
$CMD > $O/prof_${K}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_${K}_${TAG}.csv $CMD > $O/ncu_${K}_launches.log 2>&1
$CMD > $O/prof_${K}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c 1 -f -o $O/prof_${K}_${TAG} $CMD > $O/ncu_${K}_full.log 2>&1
tail -3 $O/ncu_${K}_full.log
