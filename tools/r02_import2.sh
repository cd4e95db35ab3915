#!/bin/bash
# round 2: the peer exchange's kernels on rank 0 of 2 under ncu (times + bytes), to see why the import costs 1 ms
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 --no-python tools/ncu_rank0.sh bench.py --gpus 2 \
  --workload n2v20 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-index-vs-scan --updates 0 > gpurun_out/p_import2.json 2> gpurun_out/p_import2.err
echo rc=$?; tail -3 gpurun_out/p_import2.err | cut -c1-300
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(l for l in open('gpurun_out/p_ncu_rank0.csv') if not l.startswith('==')) if len(r) > 5]
hdr = rows[0]; iid, iname, imet, ival, iunit = (hdr.index(x) for x in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault((r[iid], r[iname][:60]), {})[r[imet]] = (r[ival], r[iunit])
for (i, name), m in list(per.items())[-24:]:
    print(i, name, {k.split('.')[0][-22:]: v for k, v in m.items()})
PY
