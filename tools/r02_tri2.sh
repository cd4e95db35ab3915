#!/bin/bash
# round 2: cooperative search A/B on the LiveJournal-shape graph and scale-23, triangle blocks on/off; ncu of the MH walk kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "triangle or keyed or bulk" > gpurun_out/f_pytest.log 2>&1; tail -3 gpurun_out/f_pytest.log
run() { # name env...
  name=$1; shift
  env "$@" python bench.py --workload $WL --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-index-vs-scan --updates 2 > gpurun_out/f_${WL}_$name.json 2> gpurun_out/f_${WL}_$name.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/f_${WL}_$name.json").read().strip().splitlines()[-1])
u=d["updates"]
print("$WL $name value %.2f G ms %.2f | insert %.2f ms | rewalk %s | mh_fp %s" % (d["value"]/1e9, d["ms_per_step"], u["insert_latency_ms_median"], [round(b["insert"]["rewalk_kernel_ms"],2) for b in u["batches"]], d["parity_check"]["$WL"]["mh_fingerprint"]))
PY
}
for WL in lj n2v23; do
  run tri1_coop0 WHARF_TRI=1 WHARF_COOP_SEARCH=0
  run tri1_coop1 WHARF_TRI=1 WHARF_COOP_SEARCH=1
  run tri0_coop0 WHARF_TRI=0 WHARF_COOP_SEARCH=0
  run tri0_coop1 WHARF_TRI=0 WHARF_COOP_SEARCH=1
done
WL=n2v20
run tri1_coop1 WHARF_TRI=1 WHARF_COOP_SEARCH=1
ncu --set full --clock-control none --import-source on -k regex:walk_n2v_kernel -c 1 --launch-skip 1 -o gpurun_out/f_walk_n2v_n2v20 -f python tools/profile_pass.py n2v20 auto > gpurun_out/f_ncu.log 2>&1; tail -2 gpurun_out/f_ncu.log
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/f_traffic_n2v20.csv python tools/profile_pass.py n2v20 auto > gpurun_out/f_traffic.log 2>&1; tail -2 gpurun_out/f_traffic.log
