#!/bin/bash
# multi-GPU bench lines on an N-GPU box: tools/gpu_multi.sh N   (gpurun --gpus N -- 'bash tools/gpu_multi.sh N')
set -u
N=${1:-2}
mkdir -p gpurun_out
run() {  # $1 = tag, $2 = nproc, $3 = scan cfg, rest = env assignments
  tag=$1; np=$2; cfg=$3; shift 3
  env "$@" timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --gpus $np --steps 10 --warmup 3 --scan-cfg $cfg > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err
  echo "== $tag rc=$?"; tail -2 gpurun_out/bench_${tag}.err | cut -c1-300; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${tag}.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "stage_ms", "outputs_match_n1", "e2e_outputs_match_n1")}, "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("stage_ms"))
except Exception as e:
    print("no json:", e)
PY
}
if [ "$N" -eq 2 ]; then
  timeout 900 python -m pytest tests/test_gpu_scatter.py tests/test_gpu_parity.py -m gpu -q -k "scatter or route or csr or async or nvlink" 2>&1 | tail -3
  run n2 2 3
else
  run n8 8 3
  run n2 2 3
  run n4 4 3
  run n8_cfg4 8 4
fi
