# after the last batched-kernel change: its ncu capture, then the bench line (which then carries the fresh facts)
set -x
O=gpurun_out; TAG=r2z
timeout 600 python -m pytest tests/test_gpu_simplex.py tests/test_gpu_fuzz.py tests/test_gpu_canaries.py tests/test_cli_stream.py -m gpu -x -q 2>&1 | tail -2
bash tools/gpu/profile.sh batched $TAG > $O/profile_batched_$TAG.log 2>&1; tail -1 $O/profile_batched_$TAG.log
