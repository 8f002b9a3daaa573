timeout 600 python -m pytest tests/test_gpu_gauss_jordan.py -x -q 2>&1 | tail -2
for V in "" linear_programming_b200/variants/liblp_b200_gj_*.so; do
echo "== ${V:-default}"; LPB_LIB_PATH=$V timeout 300 python tools/bench_gj.py --reps 2 --cpu 0 2>&1 | tail -1 | cut -c1-200
done
