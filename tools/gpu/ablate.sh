# timing-only ablations of the batched kernel (variants built with LPB_EXP_* flags; answers are wrong by design)
for K in ${KERNELS:-w}; do
for V in default ${VARIANTS:-nosnap nomath nostore nostage}; do
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  echo -n "$K $V: "; LPB_BATCH_KERNEL=$K timeout 120 python tools/gpu/pivot_rate.py ${ARGS:-} 2>&1 | tail -1 | cut -c1-300
done; done
