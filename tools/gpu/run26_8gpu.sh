set -x
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29538"
timeout 300 $TR tools/gpu/sharded_parity.py 2>&1 | grep all_ranks | cut -c1-120
timeout 600 $TR tools/bench_single.py --pivots 200 --reps 2 > gpurun_out/single_${N}gpu.json 2> gpurun_out/single_${N}gpu.err; tail -c 900 gpurun_out/single_${N}gpu.json; tail -3 gpurun_out/single_${N}gpu.err
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; python -c "
import json; d=json.loads(open('gpurun_out/bench_8gpu.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_compact']['value'], d['clocks'])"; tail -3 gpurun_out/bench_${N}gpu.err
