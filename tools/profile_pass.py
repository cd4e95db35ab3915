"""One generate pass of a bench workload between cudaProfilerStart/Stop, for `ncu --profile-from-start off`:
    [WHARF_GRAPH=0] ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        --profile-from-start off --csv --log-file gpurun_out/traffic.csv python tools/profile_pass.py dw25 [coalesced|per_walker|auto] [rank world]
A warm pass runs first (allocations, graph capture), the profiled pass is the second one."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
import dynamicgraphrepresentationlearning_b200 as wharf  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "dw25"
mode = sys.argv[2] if len(sys.argv) > 2 else "auto"
rank, world = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 1)
spec = bench.WORKLOADS[workload]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
gen = {"auto": wharf.GENERATE_AUTO, "coalesced": wharf.GENERATE_COALESCED, "per_walker": wharf.GENERATE_PER_WALKER}[mode]
model = wharf.NODE2VEC if spec["model"] == "node2vec" else wharf.DEEPWALK
cfg = wharf.Config(walks_per_vertex=bench.WPV, walk_length=bench.LEN, random_walk_model=model, paramP=bench.P, paramQ=bench.Q,
                   deterministic_mode=spec["det"], seed=1, search_mode=wharf.SEARCH_SCAN, generate_mode=gen, rank=rank, world=world)
e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True)
e.generate_initial_random_walks()
torch.cuda.synchronize()
torch.cuda.profiler.start()
steps = e.generate_initial_random_walks()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print({"workload": workload, "mode": mode, "steps": steps, "ms": e.last_walk_kernel_ms(), "launches": e.kernel_launches()})
e.destroy()
