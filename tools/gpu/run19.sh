# phase timing of the shared-memory kernel at 1 and 2 CTAs per SM (are the two CTAs sweeping in phase?)
export LPB_LIB_PATH=linear_programming_b200/variants/liblp_b200_timing.so
for G in 148 296; do echo "== grid $G"; LPB_DEBUG_GRID=$G LPB_DEBUG_TIMING=1 timeout 300 python bench.py --dist D2 --batch 2048 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | grep -E "lpb timing|pivots_per_s" | tail -10 | cut -c1-200; done
