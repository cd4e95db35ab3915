# One 8-GPU box: strong-scaling sweep (tools/run_scale.sh), the driver's default command at N = 8, BASELINE config 5
# (scale-23 batch-size sweep at 8 GPUs) and config 3 (LiveJournal-shape insert stream) at 8 GPUs.
set -x
tag=${1:-m}
bash tools/run_scale.sh $tag 8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 8 > gpurun_out/${tag}_full_n8.json 2> gpurun_out/${tag}_full_n8.err; tail -2 gpurun_out/${tag}_full_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 8 --workload n2v23 --batch-sweep 1000,10000,100000,1000000 --no-e2e > gpurun_out/${tag}_n2v23_n8.json 2> gpurun_out/${tag}_n2v23_n8.err; tail -2 gpurun_out/${tag}_n2v23_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29703 bench.py --gpus 8 --workload lj --insert-only --updates 10 --no-e2e > gpurun_out/${tag}_lj_n8.json 2> gpurun_out/${tag}_lj_n8.err; tail -2 gpurun_out/${tag}_lj_n8.err
WHARF_BRANCHES=3 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29704 bench.py --gpus 8 --workload dw25 --no-e2e --updates 0 > gpurun_out/${tag}_dw25_n8_b3.json 2> gpurun_out/${tag}_dw25_n8_b3.err
python - <<EOF2
import json
d = json.loads([l for l in open("gpurun_out/${tag}_dw25_n8_b3.json") if l.startswith("{")][-1]); print("dw25 n8 branches=3", d["value"] / 1e9, d["ms_per_step"])
EOF2
