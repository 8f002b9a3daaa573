for V in "" linear_programming_b200/variants/liblp_b200_w12.so linear_programming_b200/variants/liblp_b200_w16.so; do
echo "== variant: ${V:-default(8)}"
export LPB_LIB_PATH=$V
timeout 120 python tools/gpu/dbg_batch.py 900 D2 2>&1 | tail -1 | cut -c1-160
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_tmp.json').read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d['value']/1e6,d['pivots_per_s']/1e6,d['extra']['D2_pivots_per_s']/1e6))
PY
LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep "lpb timing" | tail -17 | awk 'NR%4==1'
done
