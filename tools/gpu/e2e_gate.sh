# end-to-end rate of the fp64 host path: measured gate (default) vs lane forced on / off, same box
for MODE in default 1 0; do
  if [ $MODE = default ]; then unset LPB_HOST_NARROW; else export LPB_HOST_NARROW=$MODE; fi
  echo -n "LPB_HOST_NARROW=$MODE: "
  LPB_DEBUG_GATE=1 python bench.py --no-single --no-gj --no-cpu-baseline 2> gpurun_out/gate_$MODE.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('e2e %.4g LPs/s %.1f ms/step' % (d['e2e']['value'], d['e2e']['ms_per_step']), d['e2e']['lanes'])"
  grep gate gpurun_out/gate_$MODE.err | head -3
done
