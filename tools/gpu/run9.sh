LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep "lpb timing" | tail -9
