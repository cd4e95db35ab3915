#!/bin/bash
# round 2, first GPU call: host shape, store bandwidth, dw25 baseline of the round-1 code and its DRAM traffic per pass
mkdir -p gpurun_out
{ free -g; nproc; nvidia-smi --query-gpu=name,memory.total --format=csv; } > gpurun_out/a_host.txt 2>&1
python tools/bw_probe.py > gpurun_out/a_bw_probe.txt 2>&1
python tools/det_modes.py dw25 3 coalesced > gpurun_out/a_det_dw25.txt 2>&1
WHARF_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --profile-from-start off --csv --log-file gpurun_out/a_traffic_dw25.csv python tools/profile_pass.py dw25 coalesced > gpurun_out/a_traffic_dw25.log 2>&1
tail -3 gpurun_out/a_traffic_dw25.log
cat gpurun_out/a_host.txt gpurun_out/a_bw_probe.txt gpurun_out/a_det_dw25.txt
