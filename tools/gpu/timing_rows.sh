# per-phase cycle counters of the batched kernel (needs variants/liblp_b200_timing.so: build --variant timing -DLPB_TIMING)
export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_timing.so LPB_DEBUG_TIMING=1
for V in ${1:-w smem}; do
  for G in ${GRIDS:-296 148}; do
    echo "== variant $V grid $G"
    LPB_BATCH_KERNEL=$V LPB_DEBUG_GRID=$G timeout 300 python bench.py --no-single --no-gj --no-e2e --no-cpu-baseline --steps 1 --warmup 3 2>&1 | grep "lpb timing" | tail -${TAILN:-30}
  done
done
