python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --no-cpu-baseline --no-e2e --updates 0 > gpurun_out/s19_n1.json 2> gpurun_out/s19_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-e2e --updates 0 > gpurun_out/s19_n2.json 2> gpurun_out/s19_n2.err
python - <<EOF2
import json
for f in ("gpurun_out/s19_n1.json","gpurun_out/s19_n2.json"):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1]); print(f, d["n_gpus"], d["value"]/1e9, d["ms_per_step"])
    except Exception as ex: print(f, "failed", ex)
EOF2
