P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM ms/step %.2f each %s e2e %s"%(d["value"]/1e6,d["ms_per_step"],d["roofline"]["launch_ms_each"], d["e2e"] and round(d["e2e"]["value"]/1e6,3)))'
timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "$P"
timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "$P"
timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "$P"
timeout 600 python bench.py --no-e2e 2>/dev/null | python -c "$P"
