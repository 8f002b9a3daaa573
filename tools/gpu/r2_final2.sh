# after the last host-side change: GPU suite, bench line, batched-kernel ncu capture (its file holds the lane gate)
set -x
O=gpurun_out; TAG=r2z
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
bash tools/gpu/profile.sh batched $TAG > $O/profile_batched_$TAG.log 2>&1; tail -1 $O/profile_batched_$TAG.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -2 $O/${TAG}_bench.err
