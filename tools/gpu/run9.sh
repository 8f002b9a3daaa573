export LPB_LIB_PATH=linear_programming_b200/variants/liblp_b200_timing.so
echo "== 2 CTAs/SM"; LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep -E "lpb timing" | tail -9
echo "== 1 CTA/SM"; LPB_DEBUG_GRID=148 LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep -E "lpb timing" | tail -9
