set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -3
timeout 300 $TR tools/gpu/sharded_parity.py 2>&1 | tail -4
timeout 600 $TR tools/bench_single.py --pivots 100 --reps 2 > gpurun_out/single_2gpu.json 2> gpurun_out/single_2gpu.err; tail -c 900 gpurun_out/single_2gpu.json; tail -3 gpurun_out/single_2gpu.err
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; cut -c1-300 gpurun_out/bench_2gpu.json; python -c "
import json; d=json.loads(open('gpurun_out/bench_2gpu.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e'], d['e2e_compact'])"; tail -3 gpurun_out/bench_2gpu.err
timeout 600 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>&1 | tail -c 600
