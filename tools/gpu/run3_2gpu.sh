set -x
nvidia-smi --query-gpu=index,name --format=csv
nvidia-smi topo -m | head -12
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR tools/gpu/sharded_parity.py 2>&1 | tail -8
timeout 600 $TR tools/bench_single.py --pivots 100 --reps 2 > gpurun_out/single_2gpu.json 2> gpurun_out/single_2gpu.err; tail -c 1500 gpurun_out/single_2gpu.json; tail -5 gpurun_out/single_2gpu.err
timeout 300 python tools/bench_single.py --pivots 100 --reps 1 > gpurun_out/single_1gpu_b.json 2>&1; tail -c 1200 gpurun_out/single_1gpu_b.json
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; tail -c 2500 gpurun_out/bench_2gpu.json; tail -3 gpurun_out/bench_2gpu.err
