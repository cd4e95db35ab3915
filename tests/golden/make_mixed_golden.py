"""Batch kinds the other goldens do not mix, on the UNMODIFIED reference (oracle/_ref), walks maintained, dedup on:
deletes of edges that are mostly absent, a directed batch with self-loops, the same batch again (nothing new, walks still
re-walked), a directed delete and its re-insert.  Two graphs: RMAT scale-10 with batches that touch nearly every walk, and
RMAT scale-12 with batches of a few edges (small affected sets).  Run in the build container only.

    tests/golden/mixed_golden.json
"""
import json
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from make_golden import run_sequence  # noqa: E402
from oracle import wharf_oracle as wo  # noqa: E402

OUT = Path(__file__).resolve().parent
# kind, edges_number, batch_seed, self_loops, directed  (utility::generate_batch_of_edges arguments)
CASES = [
    {"scale": 10, "edges_half": 8000, "wpv": 5, "len": 40,
     "ops": [["del", 50, 9, 0, 0], ["ins", 200, 3, 1, 1], ["ins", 200, 3, 1, 1], ["del", 100, 4, 0, 1], ["ins", 100, 4, 0, 1]]},
    {"scale": 12, "edges_half": 30000, "wpv": 10, "len": 80,
     "ops": [["ins", 3, 1, 0, 0], ["del", 5, 2, 0, 1], ["ins", 8, 3, 1, 1], ["del", 3, 1, 0, 0], ["ins", 8, 3, 1, 1]]},
]


def main() -> None:
    assert wo.ref_available(), "build oracle/_ref first: make -C oracle ref"
    tmp = Path(tempfile.mkdtemp(prefix="wharf_mixed_"))
    for i, c in enumerate(CASES):
        n = 1 << c["scale"]
        off, adj = wo.rmat_graph(c["scale"], c["edges_half"])
        ops = [(k, wo.generate_batch(size, n, seed, bool(sl), bool(d))) for k, size, seed, sl, d in c["ops"]]
        r = run_sequence(tmp, f"mixed{i}", n, off, adj, ops, wpv=c["wpv"], len=c["len"], dedup=1, seq=1)
        c["phases"] = r["phases"]
        print(c["scale"], [ph.get("affected", ph.get("steps")) for ph in r["phases"]], [ph["m"] for ph in r["phases"]])
    gold = {"source": "oracle/_ref/wharf_ref (unmodified reference @7581ce9)", "cases": CASES}
    (OUT / "mixed_golden.json").write_text(json.dumps(gold, indent=1) + "\n")


if __name__ == "__main__":
    main()
