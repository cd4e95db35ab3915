"""Distributional golden for Metropolis-Hastings mode (deterministic_mode = false), produced by the UNMODIFIED
reference (oracle/_ref/wharf_ref, built from /root/reference by oracle/Makefile).

The reference draws from one global generator in this mode, so its corpora cannot be matched bit for bit; what can be
pinned is the LAW of the walks.  For each configuration the reference generates the corpus of the cora graph
(tests/golden/cora_graph.npz) for 20 generator seeds with one thread, and this script records, per seed, second-order
transition statistics of node2vec: over all steps (prev, cur) -> next with a position >= 1, the fraction of returns
(next == prev), of moves to a common neighbour of prev and cur, and of moves away; plus the visit share of the 64
most visited vertices.  tests/test_mh_distribution.py compares the keyed restatement (CPU oracle) and the CUDA engine
against these numbers.

    python tests/golden/make_mh_stats.py        # needs /root/reference (this container only); writes mh_stats.json
"""
import json
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import wharf_oracle as wo  # noqa: E402

OUT = Path(__file__).resolve().parent
CONFIGS = [
    # model, p, q, init
    ("node2vec", 0.5, 2.0, "weight"),
    ("node2vec", 0.5, 2.0, "random"),
    ("node2vec", 0.5, 2.0, "burnin"),
    ("node2vec", 4.0, 0.25, "weight"),
    ("deepwalk", 1.0, 1.0, "weight"),
]
SEEDS = 20
WPV, LEN = 10, 80


def transition_stats(walks: np.ndarray, offsets: np.ndarray, edges: np.ndarray) -> dict:
    """walks: [W][L] u32 (0xFFFFFFFF after a stopped walker).  Classes of every step pos -> pos + 1 with pos >= 1."""
    n = offsets.shape[0]
    row_ptr = np.append(offsets.astype(np.int64), edges.shape[0])
    keys = (np.repeat(np.arange(n, dtype=np.int64), np.diff(row_ptr)) << 32) | edges.astype(np.int64)   # sorted (src, dst)
    prev, cur, nxt = walks[:, :-2].astype(np.int64), walks[:, 1:-1].astype(np.int64), walks[:, 2:].astype(np.int64)
    ok = (nxt != 0xFFFFFFFF) & (cur != 0xFFFFFFFF) & (prev != 0xFFFFFFFF)
    prev, cur, nxt = prev[ok], cur[ok], nxt[ok]
    ret = nxt == prev
    q = (prev << 32) | nxt
    pos = np.searchsorted(keys, q)
    common = (~ret) & (pos < keys.shape[0]) & (keys[np.minimum(pos, keys.shape[0] - 1)] == q)
    total = float(prev.shape[0])
    visits = np.bincount(walks[walks != 0xFFFFFFFF].astype(np.int64), minlength=n).astype(np.float64)
    return {"steps": int(total), "return": float(ret.sum() / total), "common": float(common.sum() / total),
            "away": float(1.0 - (ret.sum() + common.sum()) / total), "visit_share": (visits / visits.sum()).tolist()}


def main():
    if not wo.ref_available():
        raise SystemExit("oracle/_ref/wharf_ref missing: run `make -C oracle ref` in the container that has /root/reference")
    z = np.load(OUT / "cora_graph.npz")
    off, adj = z["offsets"], z["edges"]
    n = int(off.shape[0])
    gold = {"source": "oracle/_ref/wharf_ref (unmodified reference), NUM_THREADS=1, cora 10x80, det=0", "graph": "cora_graph.npz",
            "wpv": WPV, "len": LEN, "configs": []}
    with tempfile.TemporaryDirectory() as tmp:
        g = Path(tmp) / "cora.graph"
        wo.write_graph(g, n, off, adj)
        for model, p, q, init in CONFIGS:
            runs, shares = [], []
            for seed in range(SEEDS):
                prefix = Path(tmp) / f"{model}_{init}_{seed}"
                wo.ref_run(["run", g, prefix, f"model={model}", f"p={p}", f"q={q}", "det=0", f"init={init}", f"wpv={WPV}", f"len={LEN}",
                            "dump=1", f"rngseed={1000 + seed}"], threads=1, timeout=600)
                st = transition_stats(wo.read_walks(f"{prefix}.walks.0.bin"), off, adj)
                shares.append(np.asarray(st.pop("visit_share")))
                runs.append(st)
            mean_share = np.mean(shares, axis=0)
            top = np.argsort(-mean_share)[:64]
            gold["configs"].append({"model": model, "p": p, "q": q, "init": init, "runs": runs, "top_vertices": top.tolist(),
                                    "top_share_mean": mean_share[top].tolist(), "top_share_std": np.std(shares, axis=0)[top].tolist()})
            print(model, p, q, init, {k: round(float(np.mean([r[k] for r in runs])), 5) for k in ("return", "common", "away")}, flush=True)
    (OUT / "mh_stats.json").write_text(json.dumps(gold, indent=1))
    print("wrote", OUT / "mh_stats.json")


if __name__ == "__main__":
    main()
