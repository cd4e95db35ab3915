"""The incremental flow of experiments/src/vertex-classification.cpp:118-197 on the UNMODIFIED reference (oracle/_ref): all
vertices loaded in isolation (WharfMH(n, m)), walks generated (every walk is its start vertex alone), then the lines of
experiments/data/cora-graph inserted as an edge stream - create_edge_stream's partitions of 1000 lines, each line as (u, v)
and (v, u), insert_edges_batch(size, edges, false, true) - with the walks maintained after every batch.  Run in the
build container only.

    tests/golden/stream_golden.json   per batch: edges, affected walks, SHA-256 of the affected ids and of the corpus
"""
import json
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from make_golden import run_sequence  # noqa: E402
from oracle import wharf_oracle as wo  # noqa: E402

OUT = Path(__file__).resolve().parent
PARTITION = 1000


def edge_stream(pairs: np.ndarray, partition: int) -> list[np.ndarray]:
    """vertex-classification.cpp:14-27 restated (the product's create_edge_stream is checked against the same loop)."""
    out = []
    for first in range(0, pairs.shape[0], partition):
        part = pairs[first: first + partition]
        out.append(np.stack([part, part[:, ::-1]], axis=1).reshape(-1, 2).astype(np.uint32))
    return out


def main() -> None:
    assert wo.ref_available(), "build oracle/_ref first: make -C oracle ref"
    pairs = np.load(OUT / "cora_edge_list.npz")["pairs"]
    n = int(pairs.max()) + 1
    stream = edge_stream(pairs, PARTITION)
    tmp = Path(tempfile.mkdtemp(prefix="wharf_stream_"))
    r = run_sequence(tmp, "stream", n, np.zeros(n, dtype=np.uint64), np.zeros(0, dtype=np.uint32), [("ins", b) for b in stream], dedup=1, seq=1)
    gold = {"source": "oracle/_ref/wharf_ref (unmodified reference @7581ce9): empty graph, then cora's lines as an edge stream",
            "n": n, "partition": PARTITION, "edge_list": "tests/golden/cora_edge_list.npz", "phases": r["phases"]}
    (OUT / "stream_golden.json").write_text(json.dumps(gold, indent=1) + "\n")
    for ph in r["phases"]:
        print({k: v for k, v in ph.items() if "sha" not in k})


if __name__ == "__main__":
    main()
