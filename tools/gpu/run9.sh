export LPB_LIB_PATH=linear_programming_b200/variants/liblp_b200_timing.so
for K in reg smem; do echo "== $K"; LPB_BATCH_KERNEL=$K LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep -E "lpb timing" | tail -17 | awk 'NR%3==1'; done
