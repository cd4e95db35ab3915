#!/bin/bash
# round 2: triangle blocks - parity suite, then n2v20 with and without them
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/e_pytest.log 2>&1
tail -12 gpurun_out/e_pytest.log
for tri in 1 0; do
  WHARF_TRI=$tri python bench.py --workload n2v20 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-index-vs-scan > gpurun_out/e_bench_n2v20_tri$tri.json 2> gpurun_out/e_bench_n2v20_tri$tri.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/e_bench_n2v20_tri$tri.json").read().strip().splitlines()[-1])
u=d["updates"]
print("tri=$tri value %.2f G ms %.2f with_index %.2f G | insert %.2f ms delete %.2f ms | graph_update %s rewalk %s | fp %s" % (d["value"]/1e9, d["ms_per_step"], (d["value_with_index"] or 0)/1e9, u["insert_latency_ms_median"], u["delete_latency_ms_median"], [round(b["insert"]["graph_update_ms"],2) for b in u["batches"]], [round(b["insert"]["rewalk_kernel_ms"],2) for b in u["batches"]], d["parity_check"]))
PY
done
