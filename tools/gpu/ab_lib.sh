# A/B of library builds on one box (variants under linear_programming_b200/variants/, "default" = the product library):
#   gpurun -- 'LIBS="default noreduce" bash tools/gpu/ab_lib.sh'
O=gpurun_out
for V in ${LIBS:-default}; do
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  for REP in 1 2; do
  timeout 300 python bench.py --no-single --no-gj --no-e2e --steps 5 > $O/abl_$V.json 2> $O/abl_$V.err || tail -3 $O/abl_$V.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/abl_$V.json") if l.startswith("{")][-1])
    print("$V", "LPs/s %.4g" % d["value"], "pivots/s %.4g" % d["pivots_per_s"], "parity", d["parity"]["bit_exact_vs_oracle"], "D2 pivots/s %.4g" % d["extra"]["D2_pivots_per_s"])
except Exception as ex:
    print("$V failed", ex)
PY
  done
done
