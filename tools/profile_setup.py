"""Engine set-up of an MH workload (layout switch, triangle blocks, first generate) under ncu's launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/setup.csv python tools/profile_setup.py n2v20"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
import dynamicgraphrepresentationlearning_b200 as wharf  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "n2v20"
spec = bench.WORKLOADS[workload]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
cfg = wharf.Config(walks_per_vertex=bench.WPV, walk_length=bench.LEN, random_walk_model=wharf.NODE2VEC, paramP=bench.P, paramQ=bench.Q,
                   deterministic_mode=False, seed=1, search_mode=wharf.SEARCH_SCAN)
torch.cuda.synchronize()
torch.cuda.profiler.start()
e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True)
steps = e.generate_initial_random_walks()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print({"workload": workload, "steps": steps, "ms": e.last_walk_kernel_ms(), "filter": e.neighbour_filter()})
e.destroy()
