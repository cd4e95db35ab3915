#!/bin/bash
# round 2: BASELINE configs 3 and 5 at 8 GPUs with the final code (updates through ShardedWharfMH: NCCL C1 + C2 inside the latency)
mkdir -p gpurun_out
run8() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 50)) bench.py --gpus 8 "$@" > gpurun_out/y_$name.json 2> gpurun_out/y_$name.err; echo "$name rc=$?"; }
run8 n2v23_n8 --workload n2v23 --batch-sweep 1000,10000,100000,1000000 --no-e2e --no-cpu-baseline --steps 5 --warmup 3
run8 lj_n8 --workload lj --insert-only --updates 10 --no-e2e --no-cpu-baseline --no-index-vs-scan --steps 5 --warmup 3
python - <<'PY'
import json
for f in ("n2v23_n8", "lj_n8"):
    d = json.loads(open(f'gpurun_out/y_{f}.json').read().strip().splitlines()[-1])
    u = d['updates']
    print(f, 'value %.1f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), 'with_index', d.get('value_with_index'))
    for r in u['batches']:
        print('   edges', r['batch_edges'], {k: {kk: (round(vv, 2) if isinstance(vv, float) else vv) for kk, vv in r[k].items() if kk in ('latency_ms', 'affected_walks', 'rewalk_kernel_ms', 'allgather_ms')} for k in ('insert', 'delete') if k in r})
    for r in (u.get('index_vs_scan') or []):
        print('   index_vs_scan', r['batch_edges'], {m: {k: round(r[m][k]['latency_ms'], 2) for k in ('insert', 'delete')} for m in ('index', 'scan') if 'unavailable' not in r[m]})
PY
