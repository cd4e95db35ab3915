set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-e2e --updates 0 > gpurun_out/g_n2v20_n2.json 2> gpurun_out/g_n2v20_n2.err; tail -2 gpurun_out/g_n2v20_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload dw25 --no-e2e --updates 0 > gpurun_out/g_dw25_n2.json 2> gpurun_out/g_dw25_n2.err; tail -2 gpurun_out/g_dw25_n2.err
for w in 4 8; do python tools/det_modes.py dw25 3 coalesced 0 $w > gpurun_out/g_det_dw25_w$w.txt 2>&1; cat gpurun_out/g_det_dw25_w$w.txt; done
python tools/det_modes.py dw25 3 coalesced 7 8 > gpurun_out/g_det_dw25_w8r7.txt 2>&1; cat gpurun_out/g_det_dw25_w8r7.txt
