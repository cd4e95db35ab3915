"""Times deterministic bulk generation in both schedules (GENERATE_PER_WALKER / GENERATE_COALESCED) on a bench workload
and checks that the corpus fingerprints agree.     python tools/det_modes.py [workload=dw25] [iters=3] [only=coalesced|per_walker]"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
import dynamicgraphrepresentationlearning_b200 as wharf  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "dw25"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
only = sys.argv[3] if len(sys.argv) > 3 else ""
rank, world = (int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (0, 1)   # time one shard of a multi-rank run          # "coalesced" / "per_walker": time one schedule only (ncu launch lists)
spec = bench.WORKLOADS[workload]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
fps = {}
for name, mode in (("per_walker", wharf.GENERATE_PER_WALKER), ("coalesced", wharf.GENERATE_COALESCED)):
    if only and name != only:
        continue
    cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.DEEPWALK, deterministic_mode=True,
                       search_mode=wharf.SEARCH_SCAN, generate_mode=mode, rank=rank, world=world)
    e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True)
    ms = []
    for it in range(iters + 1):
        flush.fill_(it)
        torch.cuda.synchronize()
        steps = e.generate_initial_random_walks()
        if it:
            ms.append(e.last_walk_kernel_ms())
    fps[name] = e.corpus_fingerprint()
    avg = sum(ms) / len(ms)
    print(json.dumps({"workload": workload, "mode": name, "rank": rank, "world": world, "ms_avg": round(avg, 3), "ms_min": round(min(ms), 3), "steps": steps,
                      "gsteps_s": round(steps / avg / 1e6, 2), "frac_of_68B_roofline": round(steps * 68 / (avg * 1e-3) / 1e9 / 6553.3, 4),
                      "launches": e.kernel_launches(), "fingerprint": fps[name]}), flush=True)
    e.destroy()
assert only or fps["per_walker"] == fps["coalesced"], "schedules disagree"
