#!/bin/bash
# One gpurun call: the whole GPU suite, then the bench line (other_configs, e2e with two sessions in flight).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -3 gpurun_out/bench.err
python - <<'P'
import json
d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step'], 'stage', d['stage_ms'])
print('e2e', json.dumps(d['e2e'])[:900])
oc = d.get('other_configs', {})
print('config5', json.dumps(oc.get('config5'))[:700])
print('config2', json.dumps(oc.get('config2'))[:500])
print('config3r', {k: v.get('total_ms') for k, v in oc.get('config3_radius', {}).items() if isinstance(v, dict)})
P
