#!/bin/bash
# round 2: BASELINE configs 3 (LiveJournal shape, stream of 100 K-insert batches) and 5 (scale-23 batch-size sweep), one GPU, final code
mkdir -p gpurun_out
timeout 900 python bench.py --workload lj --insert-only --updates 10 --no-e2e --no-cpu-baseline --no-index-vs-scan --steps 5 --warmup 3 > gpurun_out/w_lj.json 2> gpurun_out/w_lj.err
timeout 900 python bench.py --workload n2v23 --batch-sweep 1000,10000,100000,1000000 --no-e2e --no-cpu-baseline --no-index-vs-scan --steps 5 --warmup 3 > gpurun_out/w_n2v23.json 2> gpurun_out/w_n2v23.err
python - <<'PY'
import json
for f in ("lj", "n2v23"):
    d = json.loads(open(f'gpurun_out/w_{f}.json').read().strip().splitlines()[-1])
    u = d['updates']
    print(f, 'value %.2f G ms %.1f' % (d['value']/1e9, d['ms_per_step']), 'det', (d.get('deterministic_mode') or {}).get('ms_per_step'))
    for r in u['batches']:
        print('   edges', r['batch_edges'], {k: {kk: (round(vv, 1) if isinstance(vv, float) else vv) for kk, vv in r[k].items() if kk in ('latency_ms', 'affected_walks', 'rewalk_kernel_ms', 'graph_update_ms')} for k in ('insert', 'delete') if k in r})
PY
