#!/bin/bash
# One gpurun call: the radius / capped-radius parity tests, the bench line with other_configs, and the e2e pipeline trace.
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_adapters.py -m gpu -x -q -k "radius or csr or reductions or hist or stdev or capped or don or adapters" > gpurun_out/pytest_radius.log 2>&1; echo "pytest radius rc=$?"
tail -4 gpurun_out/pytest_radius.log
python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "radius" > gpurun_out/pytest_radius_full.log 2>&1; echo "pytest radius fullsize rc=$?"
tail -3 gpurun_out/pytest_radius_full.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/bench.json
CCG_PIPELINE_TRACE=1 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err; echo "trace rc=$?"
grep "ccg pipeline" gpurun_out/bench_trace.err | tail -34
