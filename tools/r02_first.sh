#!/bin/bash
# round 2: first step of the whole replicas in one sweep - parity, A/B on dw25, traffic of the new pass
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_mh_distribution.py tests/test_full_size_properties.py tests/test_walk_shapes.py -x -q -m gpu > gpurun_out/s_pytest.log 2>&1; tail -3 gpurun_out/s_pytest.log
for fs in 0 1; do
  WHARF_FIRST_STEPS=$fs timeout 900 python bench.py --workload dw25 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --updates 0 > gpurun_out/s_dw25_fs$fs.json 2> gpurun_out/s_dw25_fs$fs.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/s_dw25_fs$fs.json').read().strip().splitlines()[-1])
print('first_steps=$fs value %.1f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), d['parity_check']['dw25'])
PY
done
timeout 600 python bench.py --workload n2v20 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 2 > gpurun_out/s_n2v20.json 2> gpurun_out/s_n2v20.err
python - <<PY
import json
d = json.loads(open('gpurun_out/s_n2v20.json').read().strip().splitlines()[-1])
print('n2v20 value %.2f G ms %.2f' % (d['value']/1e9, d['ms_per_step']), 'ins', d['updates']['insert_latency_ms_median'], d['parity_check']['n2v20'].get('mh_fingerprint_equals_cpu_oracle'))
PY
bash tools/r02_setup.sh 2>&1 | grep "share_init\|tri_build\|workload"
cap() { # workload mode rank world
  out=gpurun_out/s_traffic_$1_$4.csv
  WHARF_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off --csv \
    --log-file $out python tools/profile_pass.py $1 $2 $3 $4 > gpurun_out/s_traffic_$1_$4.log 2>&1
  tail -1 gpurun_out/s_traffic_$1_$4.log | cut -c1-200
}
for world in 1 2 4; do cap dw25 coalesced 0 $world; done
