set -x
nvidia-smi --query-gpu=index,name --format=csv | head -10
for N in 8 4; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N"
timeout 300 $TR tools/gpu/sharded_parity.py 2>&1 | grep all_ranks
timeout 600 $TR tools/bench_single.py --pivots 200 --reps 2 > gpurun_out/single_${N}gpu.json 2> gpurun_out/single_${N}gpu.err; tail -c 1200 gpurun_out/single_${N}gpu.json; tail -3 gpurun_out/single_${N}gpu.err
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; cut -c1-400 gpurun_out/bench_${N}gpu.json; tail -3 gpurun_out/bench_${N}gpu.err
done
