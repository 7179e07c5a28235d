#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
CCG_KEY_BITS_LIST=16,18,21 timeout 1200 python tools/keybits_probe.py 3,4 2>&1 | tee gpurun_out/keybits.log
