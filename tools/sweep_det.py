"""GPU sweep of the bulk deterministic walk kernel's tuning knobs (WHARF_DET_VARIANT = walks,block,stream).
    python tools/sweep_det.py [workload] [variants...]      e.g.  python tools/sweep_det.py dw25 0,256,0 1,256,0 2,256,1
Prints device ms per pass and G walk-steps/s; checks that every variant yields the same corpus fingerprint."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
import dynamicgraphrepresentationlearning_b200 as wharf  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "dw25"
variants = sys.argv[2:] or ["0,256,0", "1,256,0", "1,256,1", "1,128,0", "1,512,0", "1,1024,0", "2,256,0", "2,256,1", "2,128,1", "2,512,1",
                            "4,256,0", "4,256,1", "4,128,1", "4,512,1", "8,256,1", "8,128,1"]
spec = bench.WORKLOADS[workload]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.DEEPWALK, deterministic_mode=True, search_mode=wharf.SEARCH_SCAN)
e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
ref_fp = None
out = []
for v in variants:
    os.environ["WHARF_DET_VARIANT"] = v
    ms = []
    for it in range(5):
        flush.fill_(it)
        torch.cuda.synchronize()
        steps = e.generate_initial_random_walks()
        if it >= 2:
            ms.append(e.last_walk_kernel_ms())
    fp = e.corpus_fingerprint()
    ref_fp = ref_fp or fp
    best, avg = min(ms), sum(ms) / len(ms)
    row = {"variant": v, "ms_avg": round(avg, 3), "ms_min": round(best, 3), "gsteps_s": round(steps / avg / 1e6, 2),
           "frac_of_68B_roofline": round(steps * 68 / (avg * 1e-3) / 1e9 / 6461.5, 4), "same_corpus": fp == ref_fp}
    out.append(row)
    print(json.dumps(row), flush=True)
assert all(r["same_corpus"] for r in out), "variants disagree"
