# A/B of kernel variants on ONE box.  Build each variant in-tree first, e.g.
#   cd linear_programming_b200/csrc && nvcc <flags of build.py> -DLPB_TIMING -o ../variants/liblp_b200_timing.so *.cu
# then:  gpurun -- 'VARIANTS="default timing" bash tools/gpu/ab_variants.sh'
# ("default" = the product library).  Debug knobs the library reads: LPB_BATCH_KERNEL=reg (opt-in
# register-resident variant), LPB_DEBUG_GRID=<ctas>, LPB_DEBUG_SPAD=<doubles of row padding>,
# LPB_DEBUG_TIMING=1 (needs a -DLPB_TIMING build: per-phase cycle counters on stderr).
P='import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print("D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM"%(d["value"]/1e6,d["pivots_per_s"]/1e6,d["extra"]["D2_pivots_per_s"]/1e6))'
for V in ${VARIANTS:-default}; do
  echo "== $V"
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  timeout 120 python tools/gpu/dbg_batch.py 300 D3 2>&1 | tail -1 | cut -c1-200
  timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>&1 | grep -E "lpb timing|metric" | tail -12 | python -c "
import sys
for line in sys.stdin:
    if line.startswith('{'):
        import json; d=json.loads(line); print('D1 LPs/s %.3fM pivots/s %.1fM | D2 pivots/s %.1fM'%(d['value']/1e6,d['pivots_per_s']/1e6,d['extra']['D2_pivots_per_s']/1e6))
    else: print(line.rstrip()[:160])"
done
