# Final 8-GPU check: N = 1 and N = 8 of both scaling workloads, configs 3 and 5 at 8 GPUs.
set -x
mkdir -p gpurun_out
tag=${1:-z}
python bench.py --no-e2e --updates 0 --no-cpu-baseline > gpurun_out/${tag}_n2v20_n1.json 2> gpurun_out/${tag}_n2v20_n1.err
run8() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29900 + RANDOM % 50)) bench.py --gpus 8 "$@" > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; tail -c 200 gpurun_out/${tag}_$name.err; }
run8 n2v20_n8 --no-e2e --updates 0
run8 dw25_n8 --workload dw25 --no-e2e --updates 0
run8 n2v23_n8 --workload n2v23 --batch-sweep 1000,10000,100000,1000000 --no-e2e
run8 lj_n8 --workload lj --insert-only --updates 10 --no-e2e
python - <<EOF2
import json
for f in ("n2v20_n1", "n2v20_n8", "dw25_n8", "n2v23_n8", "lj_n8"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/${tag}_{f}.json") if l.startswith("{")][-1])
        print(f, round(d["value"] / 1e9, 2), "G", round(d["ms_per_step"], 3), "ms", (d.get("updates") or {}).get("insert_latency_ms_median"))
    except Exception as ex:
        print(f, "failed", ex)
EOF2
