# compute-sanitizer over every kernel variant (run as: gpurun --timeout 2400 -- 'bash tools/gpu/sanitize.sh [tag]').
# Writes gpurun_out/sanitizer_<tool>_<tag>.txt; copy them to profiles/.  SURVEY.md section 5 "race detection".
#   memcheck   out-of-bounds / misaligned global+shared accesses
#   racecheck  shared-memory hazards between the warps of a CTA (the batched kernels, Gauss-Jordan, verify_bounds)
#   synccheck  divergent or mismatched barriers (__syncthreads, named barriers, bar.arrive)
# racecheck sees shared memory only: the LL-packet protocol of the column-sharded kernel lives in global memory and
# is covered by memcheck here and by the multi-process soak in `bench.py --gpus N`.
TAG=${1:-r2}; O=gpurun_out
CS=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
python tools/gpu/sanitize_cases.py > $O/sanitize_plain_$TAG.txt 2>&1 || { tail -20 $O/sanitize_plain_$TAG.txt; echo "cases fail WITHOUT the sanitizer"; exit 1; }
for TOOL in memcheck synccheck racecheck; do
  EXTRA=""
  [ $TOOL = racecheck ] && EXTRA="--racecheck-report all"
  timeout ${TMO:-900} $CS --tool $TOOL $EXTRA --error-exitcode 9 --print-limit 40 python tools/gpu/sanitize_cases.py ${CASES:-} > $O/sanitizer_${TOOL}_$TAG.txt 2>&1
  echo "$TOOL rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize case|Error|hazard" $O/sanitizer_${TOOL}_$TAG.txt | sort | uniq -c | sort -rn | head -30
done
