# quick N-GPU look at the column-sharded LP only (pipelined vs step-by-step loop):  N=8 bash tools/gpu/pp_multi.sh
N=${N:-8}; O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
for MODE in ${MODES:-1 0}; do
  LPB_DEBUG_SINGLE_PIPELINE=$MODE timeout 600 $TR bench.py --gpus $N --no-gj --no-e2e --no-cpu-baseline --steps 2 --soak-lps 16 > $O/ppm_$MODE.json 2> $O/ppm_$MODE.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/ppm_$MODE.json") if l.startswith("{")][-1])
    s = d["roofline_single"]; k = d["sharded_soak"]
    print("pipeline=$MODE N=$N: %.4f ms/pivot, frac %.3f, crc %d, setup %.2f ms | soak %.1f us/pivot ok=%s" % (s["ms_per_pivot"], s["frac"], s["pivot_crc"], s["setup_ms"], k["us_per_pivot"], k["every_rank_bit_exact_vs_oracle"]))
except Exception as ex:
    print("pipeline=$MODE failed:", ex); print(open("$O/ppm_$MODE.err").read()[-1500:])
PY
done
