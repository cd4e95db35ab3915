#!/bin/bash
# round 2: DRAM bytes per step of the MH walk kernel beyond the L2 (scale-23), with triangle blocks and with per-row blocks
mkdir -p gpurun_out
for tri in 1 0; do
  WHARF_TRI=$tri timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum \
    --clock-control none --profile-from-start off -k regex:"walk_n2v_kernel|n2v_share_init" --csv --log-file gpurun_out/aa_n2v23_tri$tri.csv \
    python tools/profile_pass.py n2v23 auto > gpurun_out/aa_n2v23_tri$tri.log 2>&1
  tail -1 gpurun_out/aa_n2v23_tri$tri.log | cut -c1-160
  grep -v "^==" gpurun_out/aa_n2v23_tri$tri.csv | cut -d, -f5,13- | cut -c1-200 | tail -10
done
