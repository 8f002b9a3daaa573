# Round-2 final evidence on ONE B200 (gpurun --timeout 2700 -- 'bash tools/gpu/r2_final.sh'): GPU suite, smoke, both
# bench arms, then the ncu launch lists + full captures of the three kernels, each tied to the source hash.
set -x
O=gpurun_out; TAG=${TAG:-r2z}
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference_arm.json 2> $O/ref.err; cut -c1-200 $O/${TAG}_bench_reference_arm.json
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -3 $O/${TAG}_bench.err
for K in batched single gj; do bash tools/gpu/profile.sh $K $TAG > $O/profile_${K}_$TAG.log 2>&1; tail -1 $O/profile_${K}_$TAG.log; done
python tools/gpu/verify_rate.py | tee $O/${TAG}_verify_bounds.json
