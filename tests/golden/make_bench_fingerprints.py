"""Corpus fingerprints of bench.py's CPU-sized workloads, computed by the CPU oracle (oracle/wharf_oracle.c) on the same
generator output, seed and configuration the bench uses.  bench.py prints the device fingerprints of every run next to
these (parity_check.*_equals_cpu_oracle), so a driver-run record shows corpus parity at the bench's own sizes.

    python tests/golden/make_bench_fingerprints.py [dw14 n2v20]     -> tests/golden/bench_fingerprints.json

The fingerprint is the wrapping sum over walks of the FNV-1a-64 of each walk's text form (wmh_corpus_fingerprint /
wo_corpus_fingerprint); deterministic mode is the reference's default mode (bit-exact vs the unmodified reference on
every golden fixture), MH mode is the keyed restatement.  scale-25 does not fit this container."""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
from oracle import wharf_oracle as wo  # noqa: E402

OUT = Path(__file__).with_name("bench_fingerprints.json")
names = sys.argv[1:] or ["dw14", "n2v20"]
result = json.loads(OUT.read_text()) if OUT.exists() else {}
for name in names:
    spec = bench.WORKLOADS[name]
    n = spec["n"]
    off, adj = wo.rmat_graph(spec["scale"], spec["edges_half"], bench.GRAPH_SEED)
    model = wo.NODE2VEC if spec["model"] == "node2vec" else wo.DEEPWALK
    entry = {"n": n, "m": int(adj.shape[0]), "walks_per_vertex": bench.WPV, "walk_length": bench.LEN}
    for det in ([True] if spec["det"] else [True, False]):
        t0 = time.time()
        o = wo.Oracle(n, off, adj, walks_per_vertex=bench.WPV, walk_length=bench.LEN, model=model, p=bench.P, q=bench.Q,
                      deterministic=det, init_strategy=wo.INIT_WEIGHT, seed=bench.MH_SEED)
        steps = o.generate()
        fp = wo.corpus_fingerprint(o.walks())
        entry["det_fingerprint" if det else "mh_fingerprint"] = fp
        entry["det_steps" if det else "mh_steps"] = int(steps)
        print(name, "deterministic" if det else "keyed MH", steps, fp, f"{time.time() - t0:.1f}s", flush=True)
        del o
    result[name] = entry
OUT.write_text(json.dumps(result, indent=1) + "\n")
