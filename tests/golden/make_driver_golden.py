"""What the reference's own experiment driver prints (experiments/src/throughput-latency.cpp, compiled here with the
reference's own headers, nothing of this repository involved) for cora, 10 x 80 deterministic DeepWalk, -trials 1:
the batch sizes it generated and the affected-walk counts of each insert and delete.  Run in the build container only.

    tests/golden/driver_golden.json

tests/test_unmodified_reference_callers.py holds the same driver source, linked against libwharfmh_b200.so instead, to
these numbers on the GPU, and the CPU oracle's replay of the sequence to them on the CPU."""
import json
import re
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent
FLAGS = ["-DEDGELONG", "-mcx16", "-std=c++17", "-march=native", "-Wno-subobject-linkage", "-DUSEMALLOC", "-DHOMEGROWN", "-pthread", "-O2", "-w"]


def main() -> None:
    from dynamicgraphrepresentationlearning_b200 import graph_io
    tmp = Path(tempfile.mkdtemp(prefix="wharf_driver_"))
    exe = tmp / "throughput-latency"
    inc = [f"-I{REF / d}" for d in ("libs/libcuckoo", "libs/compressed_trees", "graph", "walks", "utils", "config")]
    subprocess.run(["g++", *FLAGS, *inc, str(REF / "experiments/src/throughput-latency.cpp"), "-o", str(exe), "-ldl"], check=True)
    z = np.load(OUT / "cora_graph.npz")
    graph_io.write_unweighted_graph(tmp / "cora.adj", z["offsets"], z["edges"])
    args = ["-f", str(tmp / "cora.adj"), "-w", "10", "-l", "80", "-trials", "1"]
    out = subprocess.run([str(exe), *args], check=True, capture_output=True, text=True, env={"NUM_THREADS": "8"}).stdout
    gold = {"source": "experiments/src/throughput-latency.cpp built with the reference's own headers @7581ce9 (-O2, HOMEGROWN)",
            "args": args[2:], "graph": "tests/golden/cora_graph.npz",
            "batch_pairs": [int(x) for x in re.findall(r"Batch size = \d+ \| (\d+) ", out)],
            "walks_affected": [float(x) for x in re.findall(r"Average number of walks affected = ([0-9.e+]+)", out)]}
    (OUT / "driver_golden.json").write_text(json.dumps(gold, indent=1) + "\n")
    print(gold)


if __name__ == "__main__":
    main()
