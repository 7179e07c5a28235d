#!/bin/bash
set -u
mkdir -p gpurun_out
run() { env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) tools/bcast_probe.py 2>&1 | grep -E "^\[" | tee -a gpurun_out/bcast_probe.log; }
run NCCL_DEBUG=WARN
run NCCL_MIN_NCHANNELS=16
run NCCL_MIN_NCHANNELS=32
run NCCL_ALGO=Tree
run NCCL_ALGO=Ring NCCL_MIN_NCHANNELS=32
run NCCL_ALGO=NVLS
run NCCL_PROTO=Simple NCCL_MIN_NCHANNELS=32 NCCL_BUFFSIZE=16777216
