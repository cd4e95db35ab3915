#!/bin/bash
# round 2: the driver's command at N GPUs (default workload set), plus the 2-GPU peer-exchange test on the same box
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_exchange.py -x -q -m gpu > gpurun_out/n_peer_pytest.log 2>&1; tail -2 gpurun_out/n_peer_pytest.log
t0=$(date +%s)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 \
  > gpurun_out/n_bench_$N.json 2> gpurun_out/n_bench_$N.err
echo "rc=$? wall=$(( $(date +%s) - t0 ))s"
tail -5 gpurun_out/n_bench_$N.err | cut -c1-300
python - <<PY
import json
d = json.loads(open('gpurun_out/n_bench_$N.json').read().strip().splitlines()[-1])
print('workload', d['config']['workload'][:60], 'value %.1f G' % (d['value']/1e9), 'ms', round(d['ms_per_step'],2), 'e2e', (d.get('e2e') or {}).get('value'), (d.get('e2e') or {}).get('unavailable'))
print('roofline frac', d['roofline'].get('frac'), 'with_index', d.get('value_with_index'), d.get('index_build_ms'))
print('parity', json.dumps(d.get('parity_check'))[:900])
for name, b in d.get('blocks', {}).items():
    u = b.get('updates') or {}
    print(name, 'value %.1f G' % (b['value']/1e9), 'ms', round(b['ms_per_step'],3), 'e2e', (b.get('e2e') or {}).get('value'), 'ins', u.get('insert_latency_ms_median'), 'del', u.get('delete_latency_ms_median'), 'frac', (b.get('roofline') or {}).get('frac'))
    for r in u.get('batches', [])[:2]:
        print('   ', {k: (round(v,2) if isinstance(v,float) else v) for k,v in r['insert'].items()})
PY
