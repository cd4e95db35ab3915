#!/bin/bash
mkdir -p gpurun_out
for wl in n2v20 lj; do
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/j_setup_$wl.csv python tools/profile_setup.py $wl > gpurun_out/j_setup_$wl.log 2>&1
  tail -1 gpurun_out/j_setup_$wl.log
done
python - <<'PY'
import csv, collections
for wl in ("n2v20", "lj"):
    rows = [r for r in csv.reader(open(f"gpurun_out/j_setup_{wl}.csv")) if len(r) > 10 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        name = r[4][:70]; val = float(r[-1].replace(",", ""))
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += val
    print(wl)
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f"  {t/1e6:9.3f} ms x{c:<4d} {k}")
PY
