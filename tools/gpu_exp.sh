#!/bin/bash
# experiment call: the no-alloc test alone in a fresh process, A/B builds of the tile kernel, then the whole round (tools/gpu_round.sh)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4
timeout 600 python tools/lib_variants.py build/variants/*.so 2>&1 | tee gpurun_out/lib_variants.log
timeout 1500 bash tools/gpu_round.sh
