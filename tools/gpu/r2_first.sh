# round-2 first pass on one B200: GPU test suite, then the new one-line bench (configs 4, 5, 3)
set -x
O=gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc; free -g | sed -n 2p
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py > $O/r2a_bench.json 2> $O/r2a_bench.err; echo "bench rc=$?"; tail -c 6000 $O/r2a_bench.json; tail -5 $O/r2a_bench.err
