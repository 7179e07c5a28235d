#!/bin/bash
# one gpurun call of the kernel work loop: parity tests, A/B timings of the k_tile_sel knobs, launch list, probes
set -u
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python tools/sel_variants.py 3:1 stats 2>&1 | grep -v "^\[ccg\] tiles.*$" | tee gpurun_out/sel_variants.log
CCG_COLLECT_STATS=1 CCG_VERBOSE_STATS=1 timeout 300 python tools/prof_knn.py 1 2>&1 | tail -3 | tee gpurun_out/stats.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-other-configs"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|Device" -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
python tools/launch_times.py gpurun_out/launches.csv 2>&1 | tail -30 | tee gpurun_out/launches.txt
timeout 300 python tools/h2d_probe.py 2>&1 | tee gpurun_out/h2d_probe.log
CCG_SEL_CHUNKS=1 timeout 900 bash tools/gpu_prof.sh k_tile_sel prof_sel_r02b
