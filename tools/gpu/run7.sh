echo "== 2048 x3"; for k in 1 2 3; do timeout 120 python tools/gpu/dbg_batch.py 2048 D1 2>&1 | tail -1 | cut -c1-160; done
echo "== D2 600 x3"; for k in 1 2 3; do timeout 120 python tools/gpu/dbg_batch.py 600 D2 2>&1 | tail -1 | cut -c1-160; done
echo "== D3 600 x2"; for k in 1 2; do timeout 120 python tools/gpu/dbg_batch.py 600 D3 2>&1 | tail -1 | cut -c1-160; done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1_c.json 2> gpurun_out/bench_r1_c.err; tail -c 2600 gpurun_out/bench_r1_c.json; tail -5 gpurun_out/bench_r1_c.err
