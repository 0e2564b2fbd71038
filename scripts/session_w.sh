#!/bin/bash
# r02 session W (gpurun --gpus 8): the bench at 8, 4 and 2 ranks with the final r02 library
mkdir -p gpurun_out
run() { # n tag flags...
  n=$1; tag=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $n --steps 5 --warmup 3 "$@" 2> gpurun_out/w_bench_$tag.err | tail -1 > gpurun_out/w_bench_$tag.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/w_bench_$tag.json"))
    print("$tag", "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "ms/step %.2f" % d["ms_per_step"], "sweep_ms %.2f" % d["roofline"]["launch_ms"],
          "share %.3f" % d["roofline"]["sweep_share_of_step"], "tile", d["tile"], "parity ok", d["parity"]["ok"], "max_err %.2e" % d["parity"]["max_err"], "bit-identical ranks", d["parity"]["ranks_bit_identical"], d["clocks"])
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/w_bench_$tag.err").read()[-3000:])
PY
}
run 8 n8
run 4 n4
run 2 n2
