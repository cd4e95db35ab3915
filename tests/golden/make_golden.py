"""Generates tests/golden/* by running the UNMODIFIED reference (oracle/_ref/wharf_ref, built from
/root/reference by oracle/Makefile).  Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Outputs (committed):
    tests/golden/cora_graph.npz   cora as symmetric CSR (experiments/bin/SNAPtoAdj -s on experiments/data/cora-graph)
    tests/golden/golden.json      KATs and corpus fingerprints taken from the reference
    tests/golden/rmat8_walks.npz  full walk matrices of a tiny RMAT run (init + 3 ops)
"""
from __future__ import annotations

import hashlib
import json
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import wharf_oracle as wo  # noqa: E402

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def ref_text(*args) -> str:
    return subprocess.run([str(wo.REF_BIN), *map(str, args)], check=True, capture_output=True, text=True).stdout


def run_sequence(tmp: Path, name: str, n: int, off, adj, ops, **kw) -> dict:
    """ops = [("ins"|"del", pairs)], returns fingerprints of every phase as the reference produced them."""
    g = tmp / f"{name}.graph"
    wo.write_graph(g, n, off, adj)
    args = ["run", g, tmp / name] + [f"{k}={v}" for k, v in kw.items()]
    for i, (kind, pairs) in enumerate(ops):
        wo.write_batch(tmp / f"{name}.b{i}", pairs)
        args.append(f"{kind}:{tmp / f'{name}.b{i}'}")
    res = wo.ref_run(args)
    phases = []
    for k, r in enumerate(res):
        assert r["selfcheck_bad"] == 0, r
        walks = wo.read_walks(tmp / f"{name}.walks.{k}.bin")
        ph = {"corpus_fnv1a": wo.corpus_hash(walks), "walks_sha256": sha(walks), "m": r["m"]}
        if k == 0:
            ph["steps"] = r["steps"]
        else:
            ids = wo.read_ids(tmp / f"{name}.affected.{k}.bin")
            ph.update(kind=ops[k - 1][0], affected=int(ids.shape[0]), affected_sha256=sha(ids))
        phases.append(ph)
    return {"phases": phases, "walks": [wo.read_walks(tmp / f"{name}.walks.{k}.bin") for k in range(len(res))]}


def main() -> None:
    assert wo.ref_available(), "build oracle/_ref first: make -C oracle ref"
    gold: dict = {"source": "oracle/_ref/wharf_ref = unmodified reference @7581ce9, -O2, HOMEGROWN scheduler"}
    tmp = Path(tempfile.mkdtemp(prefix="wharf_golden_"))

    # utility::Random (utils/utility.h:152-223)
    gold["random"] = {str(s): [int(x) for x in ref_text("rng", s, 8).split()] for s in (0, 1, 2, 9, 12345, 2**63 + 5)}
    gold["drand"] = {str(s): [float(x) for x in ref_text("drand", s, 4).split()] for s in (0, 7)}

    # pairings (walks/pairings.h; tests/pairings.cpp KATs and a deterministic sample)
    rng = np.random.default_rng(1)
    p32 = [(65535, 65535), (0, 0), (1, 0), (0, 1), (123, 25)] + [tuple(int(v) for v in rng.integers(0, 65536, 2)) for _ in range(20)]
    p64 = [(2**32 - 1, 2**32 - 1), (0, 0), (5, 2**31)] + [tuple(int(v) for v in rng.integers(0, 2**32, 2)) for _ in range(20)]
    gold["pair32"] = [[x, y, int(ref_text("pair32", x, y))] for x, y in p32]
    gold["pair64"] = [[x, y, int(ref_text("pair64", x, y))] for x, y in p64]
    gold["unpair64"] = [[z] + [int(v) for v in ref_text("unpair64", z).split()] for z in
                        [0, 1, 2, 3, 99, 2**40 + 17, 2**53 - 1] + [int(v) for v in rng.integers(0, 2**52, 10)]]
    gold["triplet32"] = [[123, 25, 200, int(ref_text("triplet32", 123, 25, 200))]]
    t64 = [(123, 25, 200), (65535, 65535, 2**32 - 2)] + [(int(a), int(b), int(c)) for a, b, c in
                                                          zip(rng.integers(0, 65536, 10), rng.integers(0, 65536, 10), rng.integers(0, 2**32 - 1, 10))]
    gold["triplet64"] = [[a, b, c, int(ref_text("triplet64", a, b, c))] for a, b, c in t64]

    # generate_batch_of_edges (utils/utility.h:55-146)
    gold["genbatch"] = []
    for (e, n, s, sl, d) in [(100, 2708, 0, 0, 0), (1000, 1025, 42, 0, 1), (5000, 16385, 7, 1, 0), (131072, 16385, 42, 0, 0)]:
        ref_text("genbatch", e, n, s, sl, d, tmp / "g.b")
        b = wo.read_batch(tmp / "g.b")
        gold["genbatch"].append({"edges": e, "vertices": n, "seed": s, "self_loops": sl, "directed": d,
                                 "count": int(b.shape[0]), "first": b[0].tolist(), "last": b[-1].tolist(), "sha256": sha(b)})

    # the 6-vertex graph of tests/sampler.cpp:24-35
    off6 = np.array([0, 2, 5, 10, 13, 15], dtype=np.uint64)
    adj6 = np.array([1, 2, 0, 2, 3, 0, 1, 3, 4, 5, 1, 2, 5, 2, 5, 2, 3, 4], dtype=np.uint32)
    b6 = np.array([[0, 3], [4, 0]], dtype=np.uint32)
    r = run_sequence(tmp, "six", 6, off6, adj6, [("ins", b6), ("del", b6)], wpv=2, len=8, nn=8, dedup=1, seq=1)
    gold["six"] = {"offsets": off6.tolist(), "edges": adj6.tolist(), "batch": b6.tolist(),
                   "walks": [w.tolist() for w in r["walks"]], "phases": r["phases"]}

    # cora
    subprocess.run([str(REF / "experiments/bin/SNAPtoAdj"), "-s", "-f", str(REF / "experiments/data/cora-graph"),
                    str(tmp / "cora.adj")], check=True, capture_output=True)
    tok = (tmp / "cora.adj").read_text().split()
    n, m = int(tok[1]), int(tok[2])
    off = np.array(tok[3:3 + n], dtype=np.uint64)
    adj = np.array(tok[3 + n:3 + n + m], dtype=np.uint32)
    np.savez_compressed(OUT / "cora_graph.npz", offsets=off, edges=adj)
    b = wo.generate_batch(100, n, 0, False, False)
    r = run_sequence(tmp, "cora", n, off, adj, [("ins", b), ("del", b)], sorted=1, dedup=0, seq=1)
    gold["cora"] = {"n": n, "m": m, "batch": {"edges": 100, "vertices": n, "seed": 0}, "phases": r["phases"],
                    "walk0_prefix": r["walks"][0][0, :10].tolist(), "walk2708_prefix": r["walks"][0][2708, :10].tolist()}

    # RMAT sequences (SURVEY.md 8d recipe): both models give the same corpus in deterministic mode.
    # seq=1 (run_seq) keeps the reference's map-of-changes update race-free (graph/wharfmh.h:524-536 is a
    # check-then-act executed from parallel forks; with seq=0 a scale-14 run was seen to keep a non-minimal position).
    gold["rmat"] = []
    for scale, ehalf, model, nb, bs, wpv, L in [(8, 1500, "deepwalk", 2, 10, 3, 12), (10, 8000, "deepwalk", 4, 50, 10, 80),
                                                 (12, 30000, "node2vec", 3, 10, 10, 80), (14, 131072, "node2vec", 2, 1000, 10, 80)]:
        n = 1 << scale
        off, adj = wo.rmat_graph(scale, ehalf)
        ops, spec = [], []
        for i in range(nb):
            ops.append(("ins", wo.generate_batch(bs, n, i, False, False)))
            spec.append(["ins", bs, i])
            if i % 2 == 1:
                ops.append(("del", wo.generate_batch(bs, n, i - 1, False, False)))
                spec.append(["del", bs, i - 1])
        r = run_sequence(tmp, f"rmat{scale}", n, off, adj, ops, model=model, nn=n, wpv=wpv, len=L, seq=1)
        gold["rmat"].append({"scale": scale, "edges_half": ehalf, "graph_seed": 42, "model": model, "wpv": wpv, "len": L,
                             "ops": spec, "phases": r["phases"]})
        if scale == 8:
            np.savez_compressed(OUT / "rmat8_walks.npz", **{f"walks{k}": w for k, w in enumerate(r["walks"])})

    # MetropolisHastingsSampler distribution on the tests/sampler.cpp fixture (reference's own sampler, incl. its
    # copy-semantics: the chain's stored vertex is the init value), 20 seeds x 10000 samples per state.
    gold["sampler"] = []
    for model, p, q, cur, prev in [("deepwalk", 1, 1, 2, 2), ("node2vec", 0.7, 0.2, 2, 0), ("node2vec", 0.5, 2.0, 2, 3), ("node2vec", 0.5, 2.0, 3, 5)]:
        runs = [json.loads(ref_text("sampler", model, p, q, 10000, cur, prev, seed)) for seed in range(20)]
        gold["sampler"].append({"model": model, "p": p, "q": q, "cur": cur, "prev": prev, "iters": 10000, "runs": runs})

    (OUT / "golden.json").write_text(json.dumps(gold, indent=1))
    print("wrote", OUT / "golden.json")


if __name__ == "__main__":
    main()
