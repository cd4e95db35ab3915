#!/bin/bash
# round 2: triangle build from the longer row (set-up profile), symmetry detection tests, L2-persisting row descriptors A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_mh_distribution.py -x -q -m gpu -k "triangle or keyed or bulk or directed or chain" > gpurun_out/k_pytest.log 2>&1; tail -3 gpurun_out/k_pytest.log
bash tools/r02_setup.sh 2>&1 | grep -v "cub::\|peer_barrier\|share_splits\|repoint\|degrees_kernel"
run() { # name env...
  name=$1; shift
  env "$@" WHARF_TRACE=0 python bench.py --workload $WL --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-index-vs-scan --updates 2 > gpurun_out/k_${WL}_$name.json 2> gpurun_out/k_${WL}_$name.err
  grep "L2 persist" gpurun_out/k_${WL}_$name.err | head -1
  python - <<PY
import json
d=json.loads(open("gpurun_out/k_${WL}_$name.json").read().strip().splitlines()[-1])
u=d["updates"]
print("$WL $name value %.2f G ms %.2f | det %s | insert %.2f ms | rewalk %s | mh_fp %s" % (d["value"]/1e9, d["ms_per_step"], (d.get("deterministic_mode") or {}).get("ms_per_step"), u["insert_latency_ms_median"], [round(b["insert"]["rewalk_kernel_ms"],2) for b in u["batches"]], d["parity_check"]["$WL"]["mh_fingerprint"]))
PY
}
for WL in lj n2v23 n2v20; do
  run persist0
  run persist32 WHARF_L2_PERSIST=32
  run persist64 WHARF_L2_PERSIST=64
  run persist96 WHARF_L2_PERSIST=96
done
