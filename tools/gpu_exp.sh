#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python tools/lib_variants.py build/variants/pf4.so build/variants/pf2.so build/variants/minb14.so build/variants/finnoinl.so 2>&1 | tee gpurun_out/lib_variants.log
CCG_SEL_CHUNKS=1 SHRINK=1 bash tools/gpu_prof.sh k_tile_sel prof_sel_r02_final
