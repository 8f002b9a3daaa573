set -x
timeout 900 python -m pytest tests/test_gpu_simplex.py -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_tmp.json').read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d['value']/1e6,d['pivots_per_s']/1e6,d['extra']['D2_pivots_per_s']/1e6))
PY
tail -3 gpurun_out/bench_tmp.err
