#!/bin/bash
# r02 session Z (gpurun --gpus 2): peer mode (q read from / flux added into the owner's memory inside the sweep) - parity, then the bench against the NCCL exchange
mkdir -p gpurun_out
( timeout 400 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q -k "peer" 2>&1 | tail -5 ) | tee gpurun_out/z_pytest.txt
run() { tag=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 4 --warmup 3 "$@" 2> gpurun_out/z_$tag.err | tail -1 > gpurun_out/z_bench_$tag.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/z_bench_$tag.json"))
    print("$tag", "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "ms/step %.2f" % d["ms_per_step"], "sweep_ms %.2f" % d["roofline"]["launch_ms"],
          "share %.3f" % d["roofline"]["sweep_share_of_step"], "parity ok", d["parity"]["ok"], "max_err %.2e" % d["parity"]["max_err"], "bit-identical", d["parity"]["ranks_bit_identical"], d["exchange"][:40])
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/z_$tag.err").read()[-2500:])
PY
}
run n2_peer
run n2_nccl --no-peer
