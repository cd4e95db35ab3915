"""Extreme walk shapes on the UNMODIFIED reference (oracle/_ref): config::walks_per_vertex and config::walk_length are
uint8_t (config/globals.h:7-8), so 1 and 255 are the ends of the domain.  RMAT scale-8 (256 vertices, 1500 generated edges
per direction), an insert, the delete of the same batch, another insert, walks maintained.  Run in the build container.

    tests/golden/shape_golden.json
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
SCALE, EDGES_HALF = 8, 1500
OPS = [["ins", 20, 1], ["del", 20, 1], ["ins", 30, 2]]
SHAPES = [[1, 1], [1, 2], [2, 3], [3, 255], [255, 3], [255, 255]]


def main() -> None:
    assert wo.ref_available(), "build oracle/_ref first: make -C oracle ref"
    n = 1 << SCALE
    off, adj = wo.rmat_graph(SCALE, EDGES_HALF)
    tmp = Path(tempfile.mkdtemp(prefix="wharf_shape_"))
    cases = []
    for wpv, length in SHAPES:
        ops = [(kind, wo.generate_batch(size, n, seed, False, False)) for kind, size, seed in OPS]
        r = run_sequence(tmp, f"s{wpv}_{length}", n, off, adj, ops, wpv=wpv, len=length, dedup=1, seq=1)
        cases.append({"wpv": wpv, "len": length, "phases": r["phases"]})
        print(wpv, length, [ph.get("affected", ph.get("steps")) for ph in r["phases"]])
    gold = {"source": "oracle/_ref/wharf_ref (unmodified reference @7581ce9)", "scale": SCALE, "edges_half": EDGES_HALF, "ops": OPS, "cases": cases}
    (OUT / "shape_golden.json").write_text(json.dumps(gold, indent=1) + "\n")


if __name__ == "__main__":
    main()
