# Dantzig rule + reg-variant test + parity kernel speed check
set -x
timeout 1500 python -m pytest tests/test_gpu_simplex.py -m gpu -x -q 2>&1 | tail -8
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "$P"
timeout 300 python tools/gpu/rule_rates.py 2>&1 | tail -8
