set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_exchange.py -m gpu -x -q 2>&1 | tail -15
timeout 300 python -m pytest tests -m gpu -x -q --deselect tests/test_peer_exchange.py 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-e2e --updates 0 > gpurun_out/g_n2v20_n2.json 2> gpurun_out/g_n2v20_n2.err; tail -3 gpurun_out/g_n2v20_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --no-e2e --updates 0 --no-peers > gpurun_out/g_n2v20_n2_nopeers.json 2> gpurun_out/g_n2v20_n2_nopeers.err
timeout 300 python bench.py --no-e2e --updates 0 --no-cpu-baseline > gpurun_out/g_n2v20_n1.json 2> gpurun_out/g_n2v20_n1.err
