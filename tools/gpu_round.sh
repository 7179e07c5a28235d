#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the ncu launch list and one full capture of the top kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"^k_|Device" -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_tile_sel -s 1 -c 1 -f -o gpurun_out/prof_sel $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
# next rows: adapters program, PTX ingest timing (8.4 M lines), uncapped radius features at full size
./adapters/test_adapters > gpurun_out/adapters.log 2>&1; echo "adapters rc=$?"
python tools/ptx_bench.py 4096 2048 > gpurun_out/ptx_bench.log 2>&1; echo "ptx bench rc=$?"; tail -4 gpurun_out/ptx_bench.log
