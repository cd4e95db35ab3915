"""Generates the ingest fixtures from the reference's own tools and data (run in the build container only:
/root/reference does not exist on the GPU box, the outputs are committed).

    tests/golden/cora_edge_list.npz   experiments/data/cora-graph as a [5429, 2] array of (u, v), file order
    tests/golden/io_golden.json       sha256 of what experiments/bin/SNAPtoAdj -s writes for it (the AdjacencyGraph
                                      file whose arrays are tests/golden/cora_graph.npz), and a toy edge list with a
                                      comment, a loop and duplicates together with SNAPtoAdj -s's output for it
"""
import hashlib
import json
import subprocess
import tempfile
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent
TOY = "# comment\n3 1\n1 3\n2 2\n0 3\n0 3\n4 0\n"


def snap_to_adj(src: Path, dst: Path) -> bytes:
    subprocess.run([str(REF / "experiments/bin/SNAPtoAdj"), "-s", str(src), str(dst)], check=True, capture_output=True)
    return dst.read_bytes()


def main() -> None:
    pairs = np.loadtxt(REF / "experiments/data/cora-graph", dtype=np.uint32)
    np.savez_compressed(OUT / "cora_edge_list.npz", pairs=pairs)
    with tempfile.TemporaryDirectory() as t:
        tmp = Path(t)
        cora = snap_to_adj(REF / "experiments/data/cora-graph", tmp / "cora.adj")
        (tmp / "toy.snap").write_text(TOY)
        toy = snap_to_adj(tmp / "toy.snap", tmp / "toy.adj")
    gold = {"source": "experiments/bin/SNAPtoAdj -s of the reference @7581ce9",
            "cora_adj_sha256": hashlib.sha256(cora).hexdigest(), "cora_adj_bytes": len(cora),
            "cora_edge_list_sha256": hashlib.sha256((REF / "experiments/data/cora-graph").read_bytes()).hexdigest(),
            "toy_edge_list": TOY, "toy_adj": toy.decode()}
    (OUT / "io_golden.json").write_text(json.dumps(gold, indent=1) + "\n")
    print(gold["cora_adj_sha256"], len(cora), pairs.shape)


if __name__ == "__main__":
    main()
