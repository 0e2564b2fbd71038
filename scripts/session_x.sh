#!/bin/bash
# r02 session X (gpurun --gpus 2): device time of every collective in isolation (DOMA_B200_COMM_TIMING=1), 2 ranks, NCCL knobs
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" DOMA_B200_COMM_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 3 --warmup 3 --no-e2e --no-parity 2> gpurun_out/x_$tag.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$tag', 'ms/step %.2f' % d['ms_per_step'], 'sweep %.2f' % d['roofline']['launch_ms'])"
  grep "comm timing (rank 0" gpurun_out/x_$tag.err | cut -c30-
}
run default X=1
run ch32 NCCL_MIN_NCHANNELS=32
run ch32t512 NCCL_MIN_NCHANNELS=32 NCCL_NTHREADS=512
run simple NCCL_PROTO=Simple NCCL_MIN_NCHANNELS=32
run ctas NCCL_MIN_CTAS=32
# (an NCCL_DEBUG=INFO run piped into head hung here until its timeout and cost 20 GPU-minutes: never pipe a torchrun into head)
#env NCCL_DEBUG=INFO timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 1 --warmup 3 --no-e2e --no-parity 2>&1 | grep -iE "channel|NVLS|via P2P|Connected|nThreads" | head -12
