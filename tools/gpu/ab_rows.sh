# A/B of the batched product kernel variants on one box: parity of the forced variant first, then config 4 (D1)
# and the pivot-heavy D2 batch.   gpurun --timeout 900 -- 'bash tools/gpu/ab_rows.sh "w smem"'
O=gpurun_out
for V in ${1:-w smem}; do
  echo "== variant $V"
  LPB_BATCH_KERNEL=$V timeout 300 python -m pytest tests/test_gpu_simplex.py -m gpu -x -q -k "every_kernel_variant and $V" 2>&1 | tail -3
  LPB_BATCH_KERNEL=$V timeout 300 python bench.py --no-single --no-gj --no-e2e --steps 5 > $O/ab_$V.json 2> $O/ab_$V.err || tail -5 $O/ab_$V.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/ab_$V.json") if l.startswith("{")][-1])
    print("$V", "LPs/s %.4g" % d["value"], "pivots/s %.4g" % d["pivots_per_s"], "frac_smem %.3f" % d["roofline"]["frac"],
          "parity", d["parity"], "D2 pivots/s %.4g" % d["extra"]["D2_pivots_per_s"])
except Exception as ex:
    print("$V failed", ex)
PY
done
