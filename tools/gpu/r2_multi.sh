# N-GPU pass of the one-line bench (run as: gpurun --gpus N --timeout 1500 -- 'N=2 bash tools/gpu/r2_multi.sh')
set -x
N=${N:-2}; O=gpurun_out; TAG=${TAG:-r2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N"
nproc; free -g | sed -n 2p
[ "$N" = 2 ] && timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q -k two_processes 2>&1 | tail -3
timeout 900 $TR bench.py --gpus $N > $O/${TAG}_bench_${N}gpu.json 2> $O/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("$O/${TAG}_bench_${N}gpu.json") if l.startswith("{")][-1])
for k in ("value", "ms_per_step", "clocks", "strong_scaling", "parity", "roofline_single", "parity_single", "sharded_soak", "sharded_parity"):
    print(k, json.dumps(d.get(k))[:900])
print("e2e", d["e2e"]["value"], d["e2e"]["lanes"], "compact", d["e2e_compact"]["value"])
g = d["gauss_jordan"]; print("gj", g["ms"], g["systems_per_s"], g.get("bit_exact_vs_oracle"), g.get("checked_systems"))
PY
tail -5 $O/${TAG}_bench_${N}gpu.err
