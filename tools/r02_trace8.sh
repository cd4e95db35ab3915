#!/bin/bash
# round 2: where the per-pass fixed cost of the MH peer exchange goes at N ranks (WHARF_TRACE=1: stream synchronised at every mark)
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_exchange.py -x -q -m gpu > gpurun_out/o_peer_pytest.log 2>&1; tail -2 gpurun_out/o_peer_pytest.log
WHARF_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N \
  --workload n2v20 --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 1 > gpurun_out/o_trace_$N.json 2> gpurun_out/o_trace_$N.err
echo rc=$?
grep "wharf trace r0\]" gpurun_out/o_trace_$N.err | grep "share:\|generate: walk" | tail -24
grep "wharf trace r$((N-1))\]" gpurun_out/o_trace_$N.err | grep "share:\|generate: walk" | tail -12
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N \
  --workload n2v20 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 3 > gpurun_out/o_n2v20_$N.json 2> gpurun_out/o_n2v20_$N.err
python - <<PY
import json
d = json.loads(open('gpurun_out/o_n2v20_$N.json').read().strip().splitlines()[-1])
u = d['updates']
print('n2v20 N=$N value %.1f G ms %.3f' % (d['value']/1e9, d['ms_per_step']), 'ins', u['insert_latency_ms_median'], 'del', u['delete_latency_ms_median'])
for r in u['batches']:
    print('   ', {k: (round(v,2) if isinstance(v,float) else v) for k,v in r['insert'].items()})
PY
