#!/bin/bash
# per-kernel time/instruction list of one kNN(20)+PCA run; $1 = shrink, $2 = csv name
set -u
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"^k_|Device" --csv --log-file gpurun_out/${2:-launches}.csv python tools/prof_knn.py ${1:-2} > gpurun_out/l.log 2>&1
python tools/launch_times.py gpurun_out/${2:-launches}.csv
