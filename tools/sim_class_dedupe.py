"""CPU model of the coalesced deterministic schedule's emit gather: how many DISTINCT classes do W
consecutive walkers of a replica stand in at position p, and how much of the corpus do the K most popular classes hold?
If a warp's / block's walkers shared few classes, the record gather of coalesced_emit_kernel could be deduplicated in
shared memory; the numbers say they do not (>= 0.79 distinct per walker even for 1024-walker blocks at position 79).

    python tools/sim_class_dedupe.py 20 > profiles/r02c_class_dedupe_sim.txt"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from oracle import wharf_oracle as wo  # noqa: E402  (test infrastructure: this is an offline model, not a product path)

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = 1 << scale
off, adj = wo.rmat_graph(scale, 8 * n, seed=1)
off = off.astype(np.int64)
off = np.concatenate([off, [adj.shape[0]]]) if off.shape[0] == n else off
deg = np.diff(off)
print(f"# RMAT scale-{scale}: n {n}, directed edges {adj.shape[0]}; one replica, a fresh draw per position (the deterministic rule)")
rng = np.random.default_rng(0)
cur = np.arange(n, dtype=np.int64)
for pos in range(1, 80):
    x = int(rng.integers(0, 1 << 62))
    d = deg[cur]
    ok = d > 0
    nxt = cur.copy()
    nxt[ok] = adj[off[cur[ok]] + (x % d[ok])]
    cur = nxt
    if pos in (1, 2, 3, 4, 7, 10, 13, 16, 23, 30, 44, 58, 79):
        u, cnt = np.unique(cur, return_counts=True)
        line = f"pos {pos:2d}: classes {u.shape[0] / n:6.3f} n;"
        for blk in (32, 256, 1024):
            s = np.sort(cur.reshape(-1, blk), axis=1)
            dd = 1 + (s[:, 1:] != s[:, :-1]).sum(axis=1)
            line += f" distinct per {blk} walkers {dd.mean() / blk:5.3f};"
        cs = np.cumsum(np.sort(cnt)[::-1]) / n
        line += " top 1K/4K/16K/64K classes hold " + "/".join(f"{cs[min(k, len(cs)) - 1]:.3f}" for k in (1024, 4096, 16384, 65536))
        print(line, flush=True)
