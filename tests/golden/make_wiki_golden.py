"""A second real graph from the reference's data (experiments/data/wiki-graph: 2405 vertices, 17981 lines of which 1996
are self-loops) through the reference's own tools, for ingest and corpus parity.  Run in the build container only.

    tests/golden/wiki_edge_list.npz   the file as a [17981, 2] array of (u, v), file order
    tests/golden/wiki_graph.npz       what experiments/bin/SNAPtoAdj -s makes of it (loops and duplicates dropped)
    tests/golden/wiki_golden.json     sha256 of SNAPtoAdj's output file; corpus fingerprints of the UNMODIFIED reference
                                      (oracle/_ref) for 10 x 80 deterministic walks: initial corpus, insert of
                                      generate_batch_of_edges(200, n, 5), delete of the same, insert of (300, n, 6)
"""
import hashlib
import json
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from make_golden import run_sequence  # noqa: E402
from oracle import wharf_oracle as wo  # noqa: E402

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent


def main() -> None:
    assert wo.ref_available(), "build oracle/_ref first: make -C oracle ref"
    src = REF / "experiments/data/wiki-graph"
    pairs = np.loadtxt(src, dtype=np.uint32)
    np.savez_compressed(OUT / "wiki_edge_list.npz", pairs=pairs)
    tmp = Path(tempfile.mkdtemp(prefix="wharf_wiki_"))
    subprocess.run([str(REF / "experiments/bin/SNAPtoAdj"), "-s", str(src), str(tmp / "wiki.adj")], check=True, capture_output=True)
    text = (tmp / "wiki.adj").read_bytes()
    tok = text.split()
    n, m = int(tok[1]), int(tok[2])
    off = np.array(tok[3:3 + n], dtype=np.uint64)
    adj = np.array(tok[3 + n:3 + n + m], dtype=np.uint32)
    np.savez_compressed(OUT / "wiki_graph.npz", offsets=off, edges=adj)
    ops_spec = [["ins", 200, 5], ["del", 200, 5], ["ins", 300, 6]]
    ops = [(kind, wo.generate_batch(size, n, seed, False, False)) for kind, size, seed in ops_spec]
    r = run_sequence(tmp, "wiki", n, off, adj, ops, dedup=1, seq=1)
    gold = {"source": "experiments/bin/SNAPtoAdj -s and oracle/_ref/wharf_ref (unmodified reference @7581ce9)",
            "n": n, "m": m, "lines": int(pairs.shape[0]), "self_loops": int((pairs[:, 0] == pairs[:, 1]).sum()),
            "adj_sha256": hashlib.sha256(text).hexdigest(), "ops": ops_spec, "phases": r["phases"],
            "walk0_prefix": r["walks"][0][0, :10].tolist()}
    (OUT / "wiki_golden.json").write_text(json.dumps(gold, indent=1) + "\n")
    print(json.dumps({k: v for k, v in gold.items() if k != "phases"}))
    for ph in r["phases"]:
        print({k: v for k, v in ph.items() if "sha" not in k})


if __name__ == "__main__":
    main()
