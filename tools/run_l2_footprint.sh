# Random-gather rate against table footprint on one B200: where between 64 MiB (285 G loads/s, L2-resident) and 4 GiB
# (37.7 G, the DRAM fetch-rate wall) the cliff lies.  The MH node2vec kernel's 50 % L2 hit rate on 134 MB of uniformly
# accessed rows says the effective capacity for data every SM touches is ~67 MB (half the L2): this sweep measures it
# directly.
# usage (GPU box): bash tools/run_l2_footprint.sh <tag>
set -x
tag=${1:-l2}
mkdir -p gpurun_out tools/_build
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo tools/gather_probe.cu -o tools/_build/gather_probe || exit 1
timeout 600 tools/_build/gather_probe 64 80 96 104 112 120 128 144 160 192 256 512 > gpurun_out/${tag}_footprint.txt 2>&1
grep -A1 "^== table" gpurun_out/${tag}_footprint.txt | grep -B1 "K=1 bps=8" | grep -v "^--" | paste - - | awk '{print $3, $4, $(NF-5), $(NF-4)}' | sort -u
