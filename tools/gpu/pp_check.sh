# pipelined grid-wide loop: parity (pipeline on = default, and off), phase timing, config 5 on one GPU
set -x
SEL="tests/test_gpu_single_state.py tests/test_gpu_sharded.py tests/test_gpu_single_rules.py tests/test_gpu_canaries.py"
timeout 900 python -m pytest $SEL tests/test_gpu_simplex.py -m gpu -x -q -k "single or sharded or state or emulated or rules or golden or mid_size or grid_wide" 2>&1 | tail -4
LPB_DEBUG_SINGLE_PIPELINE=0 timeout 900 python -m pytest $SEL -m gpu -x -q -k "single or sharded or state or emulated or grid_wide" 2>&1 | tail -3
LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_timing.so timeout 300 python tools/gpu/single_phases.py 2>&1 | grep -v "^\[{" | grep -v warm -A1 | tail -12
timeout 600 python bench.py --no-gj --no-e2e --steps 2 > gpurun_out/pp_bench.json 2> gpurun_out/pp_bench.err; python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/pp_bench.json") if l.startswith("{")][-1])
s = d["roofline_single"]; print("single", s["ms_per_pivot"], s["frac"], s["pivot_crc"], s["setup_ms"]); print(d["parity_single"])
PY
tail -3 gpurun_out/pp_bench.err
