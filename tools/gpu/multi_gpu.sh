# N-GPU pass (run as: gpurun --gpus N --timeout 1500 -- 'N=8 bash tools/gpu/multi_gpu.sh'; charged N x).
# SHORT=1 skips the reference arm, REPS=1 times the 200-pivot run once.
# Sharded parity against the oracle, config 5 column-sharded, config 4 problem-sharded (both arms).
set -x
N=${N:-2}; O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N"
[ "$N" = 2 ] && timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -3
timeout 300 $TR tools/gpu/sharded_parity.py 2>&1 | grep all_ranks | cut -c1-160
timeout 600 $TR tools/bench_single.py --pivots 200 --reps ${REPS:-2} > $O/single_${N}gpu.json 2> $O/single_${N}gpu.err; tail -c 900 $O/single_${N}gpu.json; tail -3 $O/single_${N}gpu.err
timeout 600 $TR tools/bench_single.py --rule dantzig --pivots 100000 --reps 1 > $O/single_dantzig_${N}gpu.json 2> $O/single_dantzig_${N}gpu.err; tail -c 700 $O/single_dantzig_${N}gpu.json; tail -3 $O/single_dantzig_${N}gpu.err
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err; cut -c1-300 $O/bench_${N}gpu.json; grep -o '"e2e": {[^}]*}[^}]*}' $O/bench_${N}gpu.json | cut -c1-700; tail -3 $O/bench_${N}gpu.err
[ -z "$SHORT" ] && timeout 600 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>/dev/null | tail -c 500
