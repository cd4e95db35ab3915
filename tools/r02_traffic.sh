#!/bin/bash
# round 2: DRAM bytes per generate pass of one rank's shard at world sizes 2/4/8 (and the small scale-14 case), for bench.py's roofline
mkdir -p gpurun_out
cap() { # workload mode rank world
  out=gpurun_out/m_traffic_$1_$4.csv
  WHARF_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --profile-from-start off --csv \
    --log-file $out python tools/profile_pass.py $1 $2 $3 $4 > gpurun_out/m_traffic_$1_$4.log 2>&1
  tail -1 gpurun_out/m_traffic_$1_$4.log | cut -c1-200
}
for world in 2 4 8; do cap dw25 coalesced 0 $world; cap n2v20 auto 0 $world; done
cap dw14 auto 0 1
ls -la gpurun_out/m_traffic_*.csv
