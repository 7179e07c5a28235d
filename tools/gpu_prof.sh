#!/bin/bash
# one full ncu capture of the top query kernel on the config-3 scan (SHRINK=1: the full 30 M points); $1 = kernel regex, $2 = report name
set -u
mkdir -p gpurun_out
K=${1:-k_tile_sel}; NAME=${2:-prof_sel}; S=${SHRINK:-2}
python tools/prof_knn.py $S > gpurun_out/prof_plain.log 2>&1 || { tail -5 gpurun_out/prof_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -f -o gpurun_out/$NAME python tools/prof_knn.py $S > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
