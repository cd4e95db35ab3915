#!/bin/bash
# round 2: chunked inverted index (tests, scale-23 on one GPU = 6.7 G cells), triangle build after the unroll
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/l_pytest.log 2>&1; tail -3 gpurun_out/l_pytest.log
bash tools/r02_setup.sh 2>&1 | grep "tri_\|workload"
timeout 1500 python bench.py --workload n2v23 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --updates 2 > gpurun_out/l_n2v23.json 2> gpurun_out/l_n2v23.err
tail -3 gpurun_out/l_n2v23.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/l_n2v23.json').read().strip().splitlines()[-1])
print('value', d['value'] / 1e9, 'with_index', d.get('value_with_index'), 'index_build_ms', d.get('index_build_ms'), d.get('index_note', '')[:100])
print(json.dumps(d.get('index_vs_scan'), indent=None)[:2500])
PY
