#!/bin/bash
# One gpurun call: GPU parity tests, the bench line (+ reference arm), the ncu launch list and one full capture of the top kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
grep -E "passed|failed|config 5:|FAILED" gpurun_out/pytest_gpu.log | tail -12
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench.json
if [ "${REF:-0}" = "1" ]; then
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-1200 gpurun_out/bench_ref.json
fi
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-other-configs"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|Device" -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
python tools/launch_times.py gpurun_out/launches.csv > gpurun_out/launches.txt 2>&1; tail -22 gpurun_out/launches.txt
CCG_SEL_CHUNKS=1 SHRINK=1 bash tools/gpu_prof.sh k_tile_sel prof_sel_r02_final
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
./adapters/test_adapters > gpurun_out/adapters.log 2>&1; echo "adapters rc=$?"
