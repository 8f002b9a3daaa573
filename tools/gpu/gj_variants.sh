# A/B of Gauss-Jordan panel widths on one box (variants built with: python -m linear_programming_b200.build --variant gjk32 -DLPB_GJ_K=32 -DLPB_GJ_MINB=1 ;
#   --variant gjk24 -DLPB_GJ_K=24 -DLPB_GJ_MINB256=2)
V=$PWD/linear_programming_b200/variants
run() { echo -n "$1: "; shift; env "$@" python tools/bench_gj.py --reps 3 --cpu 0 --check 64 | cut -c1-260; }
run "K=16 pairs 256thr (product)" X=1
run "K=16 one-column 512thr" LPB_DEBUG_GJ_PAIRS=0
run "K=32 one-column 512thr 1 CTA/SM" LPB_LIB_PATH=$V/liblp_b200_gjk32.so LPB_DEBUG_GJ_PAIRS=0
run "K=24 one-column 256thr 2 CTA/SM" LPB_LIB_PATH=$V/liblp_b200_gjk24.so LPB_DEBUG_GJ_PAIRS=0 LPB_DEBUG_GJ_THREADS=256
