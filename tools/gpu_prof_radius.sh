#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/prof_radius.py kth 1 > gpurun_out/prof_kth_plain.log 2>&1 || { tail -5 gpurun_out/prof_kth_plain.log; exit 1; }
tail -1 gpurun_out/prof_kth_plain.log
ncu --set full --clock-control none --import-source on -k regex:k_tile_kth -s 1 -c 1 -f -o gpurun_out/r02_kth_v22 python tools/prof_radius.py kth 1 > gpurun_out/ncu_kth.log 2>&1; echo "ncu kth rc=$?"
python tools/prof_radius.py stream 1 > gpurun_out/prof_stream_plain.log 2>&1 || { tail -5 gpurun_out/prof_stream_plain.log; exit 1; }
tail -1 gpurun_out/prof_stream_plain.log
ncu --set full --clock-control none --import-source on -k regex:k_tile_stream -s 1 -c 1 -f -o gpurun_out/r02_stream_nodes_v22 python tools/prof_radius.py stream 1 > gpurun_out/ncu_stream.log 2>&1; echo "ncu stream rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|Device" -c 200 --csv --log-file gpurun_out/launches_kth.csv python tools/prof_radius.py kth 1 > /dev/null 2>&1; echo "launch list rc=$?"
ls -la gpurun_out/*.ncu-rep
