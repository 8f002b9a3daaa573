# tile height / control-group size of the pipelined loop on N GPUs (variants built with --variant rows32 -DLPB_TILE_ROWS=32 ...)
N=${N:-8}; O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N"
[ "$N" = 1 ] && TR=python
for CFG in ${CFGS:-rows32:16 rows32:8 rows16:16}; do
  V=${CFG%%:*}; C=${CFG##*:}
  if [ $V = default ]; then unset LPB_LIB_PATH; else export LPB_LIB_PATH=$PWD/linear_programming_b200/variants/liblp_b200_$V.so; fi
  LPB_DEBUG_SINGLE_NCTL=$C timeout 300 $TR bench.py --gpus $N --no-gj --no-e2e --no-cpu-baseline --steps 1 --warmup 3 --soak-lps 0 --batch 4096 > $O/ppt.json 2> $O/ppt.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("$O/ppt.json") if l.startswith("{")][-1]); s = d["roofline_single"]
    print("$CFG N=$N: %.4f ms/pivot, frac %.3f, crc %d" % (s["ms_per_pivot"], s["frac"], s["pivot_crc"]))
except Exception as ex:
    print("$CFG failed:", ex); print(open("$O/ppt.err").read()[-800:])
PY
done
