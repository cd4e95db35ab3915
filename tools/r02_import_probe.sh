#!/bin/bash
# round 2: which access of n2v_share_import_kernel costs the millisecond (2 ranks, WHARF_TRACE=1)
mkdir -p gpurun_out
for probe in 0 1 2 3; do
  WHARF_IMPORT_PROBE=$probe WHARF_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29520+probe)) bench.py --gpus 2 \
    --workload n2v20 --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 0 > gpurun_out/q_probe_$probe.json 2> gpurun_out/q_probe_$probe.err
  echo "probe $probe rc=$?"
  grep "wharf trace r0\]" gpurun_out/q_probe_$probe.err | grep "share:" | tail -4
done
