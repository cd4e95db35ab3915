#!/bin/bash
mkdir -p gpurun_out
bash tools/run_reference_tests.sh d_reftests > gpurun_out/d_reftests.log 2>&1
tail -40 gpurun_out/d_reftests.out; tail -20 gpurun_out/d_reftests.err
# destroy cost with the private pool vs the default pool
python - <<'PY' > gpurun_out/d_destroy.txt 2>&1
import time, sys, os
sys.path.insert(0, '.')
import torch, numpy as np
import bench, dynamicgraphrepresentationlearning_b200 as wharf
spec = bench.WORKLOADS["n2v20"]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=1, search_mode=wharf.SEARCH_SCAN)
for it in range(4):
    t0 = time.perf_counter(); e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True); t1 = time.perf_counter()
    e.generate_initial_random_walks(); t2 = time.perf_counter()
    e.destroy(); t3 = time.perf_counter()
    print(f"create {1e3*(t1-t0):.2f} generate {1e3*(t2-t1):.2f} destroy {1e3*(t3-t2):.2f} ms")
PY
cat gpurun_out/d_destroy.txt
WHARF_TRACE=2 python - <<'PY' > gpurun_out/d_destroy_trace.txt 2>&1
import time, sys, os
sys.path.insert(0, '.')
import torch, numpy as np
import bench, dynamicgraphrepresentationlearning_b200 as wharf
spec = bench.WORKLOADS["n2v20"]
n, m, off, adj = bench.make_graph_on_device(wharf, torch, spec, 0)
cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=1, search_mode=wharf.SEARCH_SCAN)
h = bench.pinned_empty(torch, np, (n*10, 80), np.uint32)
for it in range(3):
    t0 = time.perf_counter(); e = wharf.WharfMH(n, m, off.data_ptr(), adj.data_ptr(), config=cfg, device_arrays=True); t1 = time.perf_counter()
    e.generate_initial_random_walks_to_host(h); t2 = time.perf_counter()
    e.destroy(); t3 = time.perf_counter()
    print(f"create {1e3*(t1-t0):.2f} generate_to_host {1e3*(t2-t1):.2f} destroy {1e3*(t3-t2):.2f} ms", flush=True)
PY
tail -30 gpurun_out/d_destroy_trace.txt
