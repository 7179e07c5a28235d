#!/bin/bash
# quick developer loop on the GPU box: parity tests, then timings (stats off), then leaf statistics (stats on)
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python tools/quick_gpu.py ${1:-1:1 3:2 3:1 5:1} 2>&1 | tee gpurun_out/quick.log
CCG_COLLECT_STATS=1 CCG_VERBOSE_STATS=1 python tools/quick_gpu.py 3:2 2>&1 | tee gpurun_out/quick_stats.log
