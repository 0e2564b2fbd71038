#!/bin/bash
# r02 session D (gpurun --gpus 2): the multi-rank paths on real GPUs - cell-sharded and replicated elementwise phases
mkdir -p gpurun_out
nvidia-smi -L
( timeout 1200 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_driver.py -m gpu -q -k "two" 2>&1 | tail -15 ) | tee gpurun_out/d_pytest.txt
for mode in "" "--no-cell-sharding"; do
  tag=${mode:+replicated}; tag=${tag:-sharded}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 3 --warmup 3 $mode 2> gpurun_out/d_bench_n2_$tag.err | tail -1 > gpurun_out/d_bench_n2_$tag.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/d_bench_n2_$tag.json"))
    print("$tag", "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "ms/step %.1f" % d["ms_per_step"], "sweep_ms %.2f" % d["roofline"]["launch_ms"],
          "share %.3f" % d["roofline"]["sweep_share_of_step"], "parity", d["parity"])
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/d_bench_n2_$tag.err").read()[-3000:])
PY
done
# flat-builder start-up of the drop-in driver on the bench case (host-side; setup_seconds is the mesh construction)
for f in --flat --no-flat; do
  echo "== doma_b200_driver -i full_syn256_b.xml $f"
  timeout 900 doma_b200/bin/doma_b200_driver -i tests/golden/cases/full_syn256_b.xml $f 2>&1 | grep -E "^DRIVER" | sed 's/ k=[^ ]*//'
done 2>&1 | tee gpurun_out/d_driver_flat.txt
