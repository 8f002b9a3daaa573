# row-stride padding sweep: bank alignment of warps that straddle two 51-lane row groups
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
for PAD in 0 2 4 6 8 10 12 14 16; do
  echo "== SPAD $PAD"
  LPB_DEBUG_SPAD=$PAD timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 3 2>/dev/null | python -c "$P"
done
LPB_DEBUG_SPAD=8 timeout 120 python tools/gpu/dbg_batch.py 300 D3 2>&1 | tail -1 | cut -c1-200
