#!/bin/bash
# torchrun worker wrapper: rank 0 runs under ncu (single-pass metrics only: kernels that wait for peers must not be replayed)
if [ "$RANK" = "0" ]; then
  exec ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum \
    --clock-control none --csv --log-file gpurun_out/p_ncu_rank0.csv -k regex:"share|import|barrier|walk_n2v|invalidate" python "$@"
else
  exec python "$@"
fi
