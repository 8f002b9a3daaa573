export LPB_BATCH_KERNEL=reg
timeout 120 python tools/gpu/dbg_batch.py 96 D1 2>&1 | tail -3 | cut -c1-300
timeout 120 python tools/gpu/dbg_batch.py 600 D2 2>&1 | tail -2 | cut -c1-300
timeout 120 python tools/gpu/dbg_batch.py 600 D3 2>&1 | tail -2 | cut -c1-300
unset LPB_BATCH_KERNEL
timeout 900 python -m pytest tests/test_gpu_simplex.py tests/test_gpu_fuzz.py -x -q 2>&1 | tail -4
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
for K in reg smem; do echo "== $K"; LPB_BATCH_KERNEL=$K timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "$P"; done
