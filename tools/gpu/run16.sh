# setup-code placement experiment: noinline (default lib) vs forceinline variant, same box
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
for rep in 1 2; do
for V in default inline; do
  echo "== $V"
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "$P"
done
done
unset LPB_LIB_PATH
timeout 900 python -m pytest tests/test_gpu_simplex.py -m gpu -x -q 2>&1 | tail -2
