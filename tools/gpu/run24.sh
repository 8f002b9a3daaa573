set -x
python tools/gpu/latency_probe.py
timeout 900 python tools/bench_course.py > gpurun_out/course.json 2> gpurun_out/course.err; python - <<'PY'
import json
d=json.load(open("gpurun_out/course.json"))
for r in d["config_1_course_instances"][:6]: print({k:(round(v,1) if isinstance(v,float) else v) for k,v in r.items()})
print(d["config_2_stress"])
PY
tail -3 gpurun_out/course.err
