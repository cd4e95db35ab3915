set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/g_full_n2.json 2> gpurun_out/g_full_n2.err; tail -3 gpurun_out/g_full_n2.err
timeout 600 python tools/det_modes.py dw25 3 coalesced > gpurun_out/g_det_dw25.txt 2>&1; cat gpurun_out/g_det_dw25.txt
