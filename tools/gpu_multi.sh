#!/bin/bash
# multi-GPU bench lines on an N-GPU box: tools/gpu_multi.sh N [extra bench args]   (gpurun --gpus N -- 'bash tools/gpu_multi.sh N')
set -u
N=${1:-2}; shift || true
mkdir -p gpurun_out
run() {  # $1 = tag, $2 = nproc, rest = env assignments
  tag=$1; np=$2; shift 2
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
      bench.py --gpus $np --steps 10 --warmup 3 "${EXTRA[@]}" > gpurun_out/bench_${tag}.json 2> gpurun_out/bench_${tag}.err
  echo "== $tag rc=$?"; tail -2 gpurun_out/bench_${tag}.err | cut -c1-300; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${tag}.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "ms_per_step", "stage_ms", "outputs_match_n1", "e2e_outputs_match_n1")}, d["e2e"]["value"], d["e2e"]["ms_per_step"])
except Exception as e:
    print("no json:", e)
PY
}
EXTRA=("$@")
if [ "$N" -ge 2 ]; then timeout 600 python -m pytest tests/test_gpu_scatter.py -m gpu -q 2>&1 | tail -3; fi
np=2
while [ $np -le $N ]; do
  run n${np}_pull $np CCG_BENCH_REPL=pull
  run n${np}_bcast $np CCG_BENCH_REPL=bcast
  np=$((np * 2))
done
