#!/bin/bash
# round 2: triangle-block maintenance handling each new undirected edge once - parity and the insert batch's graph update time
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_mh_distribution.py -x -q -m gpu > gpurun_out/t_pytest.log 2>&1; tail -2 gpurun_out/t_pytest.log
timeout 600 python bench.py --workload n2v20 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 4 > gpurun_out/t_n2v20.json 2> gpurun_out/t_n2v20.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/t_n2v20.json').read().strip().splitlines()[-1])
u = d['updates']
print('n2v20 value %.2f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), 'ins', u['insert_latency_ms_median'], 'del', u['delete_latency_ms_median'])
print('graph_update insert', [round(b['insert']['graph_update_ms'], 2) for b in u['batches']], 'delete', [round(b['delete']['graph_update_ms'], 2) for b in u['batches']])
PY
for wl in lj; do
timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 2 > gpurun_out/t_$wl.json 2> gpurun_out/t_$wl.err
python - <<PY
import json
d = json.loads(open('gpurun_out/t_$wl.json').read().strip().splitlines()[-1])
u = d['updates']
print('$wl value %.2f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), 'ins', u['insert_latency_ms_median'], 'graph_update insert', [round(b['insert']['graph_update_ms'], 2) for b in u['batches']], d['parity_check']['$wl'].get('mh_fingerprint'))
PY
done
