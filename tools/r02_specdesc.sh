#!/bin/bash
# round 2: descriptor of the next vertex loaded after the accept decision (WHARF_SPEC_DESC=1 = before it, as in round 1): parity, A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_mh_distribution.py -x -q -m gpu -k "keyed or triangle or bulk or mh or chain or to_host or directed" > gpurun_out/sd_pytest.log 2>&1; tail -2 gpurun_out/sd_pytest.log
for WL in n2v20 lj n2v23; do
for spec in 1 0; do
  WHARF_SPEC_DESC=$spec timeout 600 python bench.py --workload $WL --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 2 > gpurun_out/sd_${WL}_$spec.json 2> gpurun_out/sd_${WL}_$spec.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/sd_${WL}_$spec.json').read().strip().splitlines()[-1])
u = d['updates']
print('$WL spec_desc=$spec value %.2f G ms %.2f | insert %.2f ms | rewalk %s | fp %s %s' % (d['value']/1e9, d['ms_per_step'], u['insert_latency_ms_median'], [round(b['insert']['rewalk_kernel_ms'], 2) for b in u['batches']], d['parity_check']['$WL']['mh_fingerprint'], d['parity_check']['$WL'].get('mh_fingerprint_equals_cpu_oracle')))
PY
done; done
