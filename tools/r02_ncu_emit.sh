#!/bin/bash
# round 2: smoke() of the final tree, then `ncu --set full` of the two kernels that carry the scale-25 pass
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/u_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/u_smoke.log
WHARF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"coalesced_emit_kernel" --launch-skip 60 -c 1 \
  -o gpurun_out/u_emit_dw25 -f python tools/profile_pass.py dw25 coalesced > gpurun_out/u_ncu_emit.log 2>&1; tail -1 gpurun_out/u_ncu_emit.log | cut -c1-200
WHARF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"coalesced_advance_kernel" --launch-skip 1 -c 1 \
  -o gpurun_out/u_advance_dw25 -f python tools/profile_pass.py dw25 coalesced > gpurun_out/u_ncu_adv.log 2>&1; tail -1 gpurun_out/u_ncu_adv.log | cut -c1-200
WHARF_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"coalesced_first_steps_kernel" -c 1 \
  -o gpurun_out/u_first_dw25 -f python tools/profile_pass.py dw25 coalesced > gpurun_out/u_ncu_first.log 2>&1; tail -1 gpurun_out/u_ncu_first.log | cut -c1-200
ls -la gpurun_out/u_*.ncu-rep
