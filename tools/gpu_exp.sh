#!/bin/bash
# one gpurun call of the kernel work loop: parity tests, A/B timings of the k_tile_sel knobs, stats, ncu capture
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python tools/sel_variants.py 3:1 2>&1 | tee gpurun_out/sel_variants.log
CCG_COLLECT_STATS=1 CCG_VERBOSE_STATS=1 timeout 300 python tools/prof_knn.py 1 2>&1 | tail -3 | tee gpurun_out/stats.log
CCG_SEL_CHUNKS=1 timeout 900 bash tools/gpu_prof.sh k_tile_sel prof_sel_r02b
