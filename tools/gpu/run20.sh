# sweep v2 (loads before the f test, stride pinned in a register) vs v1, same box
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
for V in $VARIANTS; do
  echo "== $V"
  export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so
  timeout 120 python tools/gpu/dbg_batch.py 300 D3 2>&1 | tail -1 | cut -c1-200
  timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "$P"
done
