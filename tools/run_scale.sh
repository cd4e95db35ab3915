# Strong-scaling sweep on one box: bench.py at N = 1, 2, 4, 8 for the headline workload (n2v20) and the scale-25 DeepWalk one.
# usage: bash tools/run_scale.sh <tag> [max_gpus]
set -x
tag=${1:-s}
maxg=${2:-8}
mkdir -p gpurun_out
port=29600
for wl in n2v20 dw25; do
  for n in 1 2 4 8; do
    [ $n -gt $maxg ] && continue
    port=$((port+1))
    if [ $n -eq 1 ]; then
      timeout 600 python bench.py --workload $wl --no-e2e --updates 0 --no-cpu-baseline > gpurun_out/${tag}_${wl}_n$n.json 2> gpurun_out/${tag}_${wl}_n$n.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --workload $wl --no-e2e --updates 0 > gpurun_out/${tag}_${wl}_n$n.json 2> gpurun_out/${tag}_${wl}_n$n.err
    fi
    tail -c 300 gpurun_out/${tag}_${wl}_n$n.err
  done
done
python - <<EOF2
import json, glob
for wl in ("n2v20", "dw25"):
    base = None
    for n in (1, 2, 4, 8):
        try:
            d = json.loads([l for l in open(f"gpurun_out/${tag}_{wl}_n{n}.json") if l.startswith("{")][-1])
        except Exception as ex:
            print(wl, n, "failed", ex); continue
        base = base or d["value"]
        print(wl, n, round(d["value"] / 1e9, 2), "G steps/s", round(d["ms_per_step"], 3), "ms", "x%.2f" % (d["value"] / base), d.get("deterministic_mode", {}).get("value"))
EOF2
