# Same-box A/B of the once-per-LP reader of A (LPB_LOAD_DEPTH: fp64 path, LPB_LOAD_DEPTH_TYPED: int8..float32 path).
# Build the variants here first (nvcc cross-compiles), e.g.
#   for d in 1 2 8; do nvcc <flags of build.py> -DLPB_LOAD_DEPTH_TYPED=$d -o linear_programming_b200/variants/liblp_b200_t$d.so linear_programming_b200/csrc/*.cu linear_programming_b200/csrc/host_narrow.cpp; done
# then:  gpurun -- 'VARIANTS="default t1 t2 t8" bash tools/gpu/ab_load_depth.sh'
# Prints `value` (fp64 A resident in HBM) and `e2e_compact` (int8 A from pinned host memory) per variant.
for V in ${VARIANTS:-default}; do
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  echo -n "$V: "
  python bench.py --no-cpu-baseline --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('value %.4f M  e2e %.4f M  e2e_compact %.4f M LPs/s  launches %s' % (d['value']/1e6, d['e2e']['value']/1e6, d['e2e_compact']['value']/1e6, d['roofline']['launch_ms_each']))"
done
