#!/bin/bash
# round 2: emit<8> variants (walkers per thread x resident blocks) on dw25
mkdir -p gpurun_out
for v in 0 1 2 3; do
  WHARF_EMIT_VARIANT=$v timeout 600 python bench.py --workload dw25 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --updates 0 > gpurun_out/v_dw25_$v.json 2> gpurun_out/v_dw25_$v.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/v_dw25_$v.json').read().strip().splitlines()[-1])
print('emit variant $v: value %.1f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), d['parity_check']['dw25']['det_fingerprint'], d['parity_check']['dw25']['schedules_equal'])
PY
done
