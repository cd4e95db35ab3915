set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-e2e --updates 0 > gpurun_out/g_n2v20_n2.json 2> gpurun_out/g_n2v20_n2.err; tail -2 gpurun_out/g_n2v20_n2.err
timeout 600 python bench.py --workload lj --insert-only --updates 10 --no-e2e --no-cpu-baseline > gpurun_out/g_lj_n1.json 2> gpurun_out/g_lj_n1.err; tail -2 gpurun_out/g_lj_n1.err
timeout 600 python bench.py --workload n2v23 --batch-sweep 1000,10000,100000,1000000 --no-e2e --no-cpu-baseline > gpurun_out/g_n2v23_n1.json 2> gpurun_out/g_n2v23_n1.err; tail -2 gpurun_out/g_n2v23_n1.err
