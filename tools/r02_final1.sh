#!/bin/bash
# round 2: one-GPU validation of the final code: the GPU suite, smoke(), the driver's two bench commands, the launch list
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r_pytest.log 2>&1; tail -3 gpurun_out/r_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r_smoke.log
t0=$(date +%s)
timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r_bench_1.json 2> gpurun_out/r_bench_1.err; echo "bench rc=$? wall=$(( $(date +%s) - t0 ))s"
t0=$(date +%s)
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r_bench_ref.json 2> gpurun_out/r_bench_ref.err; echo "reference rc=$? wall=$(( $(date +%s) - t0 ))s"
tail -c 600 gpurun_out/r_bench_ref.json
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r_bench_1.json').read().strip().splitlines()[-1])
print('workload', d['config']['workload'][:60], 'value %.1f G' % (d['value']/1e9), 'ms', round(d['ms_per_step'],2), 'e2e', (d.get('e2e') or {}).get('value'), (d.get('e2e') or {}).get('ms_per_step'))
print('roofline frac', d['roofline'].get('frac'), 'with_index', d.get('value_with_index'), 'cpu', d['cpu_baseline'].get('value'), d['cpu_baseline'].get('sample','')[:80])
print('clocks', d['clocks'])
for name, b in d.get('blocks', {}).items():
    u = b.get('updates') or {}
    print(name, 'value %.1f G' % (b['value']/1e9), 'ms', round(b['ms_per_step'],3), 'with_index', b.get('value_with_index'), 'e2e', (b.get('e2e') or {}).get('value'), (b.get('e2e') or {}).get('ms_per_step'), 'ins', u.get('insert_latency_ms_median'), 'del', u.get('delete_latency_ms_median'), 'frac', (b.get('roofline') or {}).get('frac'))
PY
echo "launch list: see profiles/r02g_launches_default_1gpu.txt"
