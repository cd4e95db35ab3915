set -x
mkdir -p gpurun_out
tag=${1:-n}
port=29800
for n in 1 2 4 8; do
  port=$((port+1))
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --workload dw25 --no-e2e --updates 0 --no-cpu-baseline > gpurun_out/${tag}_dw25_n$n.json 2> gpurun_out/${tag}_dw25_n$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --workload dw25 --no-e2e --updates 0 > gpurun_out/${tag}_dw25_n$n.json 2> gpurun_out/${tag}_dw25_n$n.err
  fi
  tail -c 200 gpurun_out/${tag}_dw25_n$n.err
done
python - <<EOF2
import json
base = None
for n in (1, 2, 4, 8):
    d = json.loads([l for l in open(f"gpurun_out/${tag}_dw25_n{n}.json") if l.startswith("{")][-1]); base = base or d["value"]
    print("dw25", n, round(d["value"] / 1e9, 1), "G steps/s", round(d["ms_per_step"], 3), "ms x%.2f" % (d["value"] / base))
EOF2
