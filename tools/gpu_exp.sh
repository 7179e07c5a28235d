#!/bin/bash
# one gpurun call of the kernel work loop: parity tests, then the A/B timings of the k_tile_sel knobs
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python tools/sel_variants.py 3:1 stats 2>&1 | tee gpurun_out/sel_variants.log
