#!/bin/bash
# multi-GPU bench on one box: N=$1 ranks, tiles from $TILES; prints one summary line per run and keeps the JSON lines
N=${1:-2}; shift
mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -12 > gpurun_out/topo_n$N.txt
for tile in ${TILES:-auto}; do
  if [ "$tile" = auto ]; then TF=""; else TF="--tile $tile"; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps ${STEPS:-3} --warmup 3 $TF "$@" 2> gpurun_out/mg_n${N}_$tile.err | tail -1 > gpurun_out/mg_n${N}_$tile.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/mg_n${N}_$tile.json"))
    print("N", d["n_gpus"], "tile", d["tile"], "value %.4g" % d["value"], "e2e %.4g" % (d["e2e"] or {}).get("value", 0), "sweep_ms %.2f" % d["roofline"]["launch_ms"],
          "ms_per_step %.1f" % d["ms_per_step"], "share %.3f" % d["roofline"]["sweep_share_of_step"], "units/rank", d["units_per_rank"], d["clocks"])
except Exception as e:
    print("run failed:", e); print(open("gpurun_out/mg_n${N}_$tile.err").read()[-3000:])
PY
done
