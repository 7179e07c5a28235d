#!/bin/bash
# 8-GPU box session: config 3 at N = 8 (both replication modes) and N = 4, config 4 (100 M points) at N = 8
set -u
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
run n8_pull_staged 8 3 CCG_BENCH_REPL=pull CCG_BENCH_OUT=staged
run n8_pull_route 8 3 CCG_BENCH_REPL=pull CCG_BENCH_OUT=route
run n8_cfg4_staged 8 4 CCG_BENCH_REPL=pull CCG_BENCH_OUT=staged
run n8_cfg4_route 8 4 CCG_BENCH_REPL=pull CCG_BENCH_OUT=route
