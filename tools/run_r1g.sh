set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/peer_probe.py > gpurun_out/g_peer_probe.txt 2>&1; grep -v Traceback gpurun_out/g_peer_probe.txt | grep "^\[" | head -20
python tools/det_modes.py dw25 3 > gpurun_out/g_det_dw25.txt 2>&1; cat gpurun_out/g_det_dw25.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload dw25 --no-e2e --updates 0 > gpurun_out/g_dw25_n2.json 2> gpurun_out/g_dw25_n2.err
tail -3 gpurun_out/g_dw25_n2.err
