"""The drop-in claim, checked with the reference's own callers (SURVEY 8b "Callers of the boundary").

include/compat/ holds forwarding headers with the reference's header names (wharfmh.h, pairings.h, pbbslib/...), so the
reference's experiment drivers and test files compile UNCHANGED against include/wharfmh_b200.hpp and link
libwharfmh_b200.so - no line of them is edited or copied; they are compiled where they lie under /root/reference
(tests/cpp/build_reference_callers.py, also run by __graft_entry__.build()).  The binaries land in tests/cpp/_build/
(git-ignored, but they travel to the GPU box like the library):

  * here (no GPU): all four experiments/src drivers and tests/wharfmh.cpp compile and link; tests/pairings.cpp also RUNS -
    its known-answer and 10 M round-trip tests pass against the library's pairing functions;
  * on the B200: the reference's throughput-latency driver runs on cora and the batch sizes and affected-walk counts it
    prints equal what the same source prints when built with the reference's own headers
    (tests/golden/driver_golden.json) and the CPU oracle's replay of the sequence; memory-footprint runs to completion.
googletest is not installed: tests/cpp/mini_gtest/ is a 60-line runner with its spelling (test infrastructure)."""
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tests" / "cpp"))
import build_reference_callers as brc  # noqa: E402

needs_reference = pytest.mark.skipif(not brc.reference_present(), reason="/root/reference is not on this machine")


@needs_reference
def test_reference_drivers_and_tests_compile_unchanged():
    built = brc.build_all()
    assert {p.name for p in built} == {"throughput-latency", "memory-throughput-latency", "memory-footprint", "vertex-classification",
                                       "ref_test_wharfmh", "ref_test_pairings", "intree_stub"}
    assert all(p.exists() for p in built)
    # nothing of the reference was copied next to the binaries or into the compat headers
    assert not list((ROOT / "tests" / "cpp" / "_build" / "ref_callers").glob("*.cpp"))
    for header in (ROOT / "include" / "compat").rglob("*.h"):
        assert len(header.read_text().splitlines()) < 60, header


@needs_reference
def test_reference_pairing_tests_pass_against_the_library():
    brc.build_all()
    r = subprocess.run([str(brc.OUT / "ref_test_pairings")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "[  TOTAL   ] 4 tests, 0 failures" in r.stdout
    assert "Encoded value (Szudzik pairing): 4294967295" in r.stdout and "Decoded value (Szudzik pairing): 65535 65535" in r.stdout   # tests/pairings.cpp:29-37
    assert "Encoded value (Szudzik pairing for triplets): 229643916" in r.stdout                                                      # :75-90
    assert "Decoded value (Szudzik pairing for triplets): 123 25 200" in r.stdout
    assert "failed" not in r.stderr                     # the 10 M round trips report mismatches on stderr (:61-64, :113-117)


def test_drivers_fail_loudly_without_gpu(tmp_path):
    exe = brc.OUT / "memory-footprint"
    if not exe.exists():
        pytest.skip("reference callers not built on this machine")
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from dynamicgraphrepresentationlearning_b200 import graph_io
    z = np.load(ROOT / "tests" / "golden" / "cora_graph.npz")
    graph_io.write_unweighted_graph(tmp_path / "cora.adj", z["offsets"], z["edges"])
    r = subprocess.run([str(exe), "-f", str(tmp_path / "cora.adj")], capture_output=True, text=True)
    assert "Vertices: 2708 Edges: 10556" in r.stdout                    # the reader ran (host code) ...
    assert r.returncode != 0 and "no CPU path" in r.stderr              # ... the engine refused to


def test_reference_and_library_in_one_binary_without_gpu(golden):
    """tests/cpp/intree_stub.cpp = INTEGRATION.md section 2 made real: the reference's own graph/wharfmh.h and the C ABI in one
    translation unit.  Without a GPU the reference half runs (its corpus equals the golden) and wmh_create refuses."""
    exe = brc.OUT / "intree_stub"
    if not exe.exists():
        pytest.skip("reference callers not built on this machine")
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 3 and "no CPU path" in r.stderr
    got = [l.split(": ", 1)[1] for l in r.stdout.splitlines() if l.startswith("reference ")]
    assert got == ["".join(f"{v} " for v in w) for w in golden["six"]["walks"][0]]


@pytest.mark.gpu
def test_reference_and_library_side_by_side_on_gpu(golden, cora, tmp_path):
    """The same binary on the B200: the UNMODIFIED reference (host threads) and the CUDA engine hold the same graph, take
    the same insert and delete, and every walk of both is compared as text - first the six-vertex fixture, then cora
    (27 080 walks, the reference's own generator for the batch)."""
    exe = brc.OUT / "intree_stub"
    if not exe.exists():
        pytest.skip("reference callers were not built where the snapshot was taken")
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "initial: 12 walks, differing 0, edges 18 / 18" in r.stdout
    assert "inserted: affected 12 / 12" in r.stdout and "inserted: 12 walks, differing 0, edges 20 / 20" in r.stdout
    assert "deleted: affected 12 / 12" in r.stdout and "deleted: 12 walks, differing 0, edges 18 / 18" in r.stdout
    assert "total differing: 0" in r.stdout
    from dynamicgraphrepresentationlearning_b200 import graph_io
    n, off, adj = cora
    graph_io.write_unweighted_graph(tmp_path / "cora.adj", off, adj)
    r = subprocess.run([str(exe), str(tmp_path / "cora.adj"), "100"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    g = golden["cora"]["phases"]
    assert f"initial: {10 * n} walks, differing 0, edges {g[0]['m']} / {g[0]['m']}" in r.stdout
    assert f"inserted: affected {g[1]['affected']} / {g[1]['affected']}" in r.stdout and f"differing 0, edges {g[1]['m']} / {g[1]['m']}" in r.stdout
    assert f"deleted: affected {g[2]['affected']} / {g[2]['affected']}" in r.stdout and "total differing: 0" in r.stdout


def replay_throughput_latency(wo, n, off, adj):
    """The driver's sequence on the CPU oracle: generate, then per batch size (throughput-latency.cpp:87-90) an insert and a
    delete of generate_batch_of_edges(size, n, 0, false, true)."""
    o = wo.Oracle(n, off, adj, walks_per_vertex=10, walk_length=80)
    o.generate()
    counts, sizes = [], []
    for size in (5, 50, 500):
        b = wo.generate_batch(size, n, 0, False, True)
        sizes.append(int(b.shape[0]))
        for delete in (False, True):
            o.update(b.copy(), delete=delete, remove_dups=True)
            counts.append(float(o.affected()[0].shape[0]))
    o.close()
    return sizes, counts


def driver_golden():
    import json
    return json.loads((ROOT / "tests" / "golden" / "driver_golden.json").read_text())


def test_oracle_replays_what_the_reference_driver_prints(wo, cora):
    """tests/golden/driver_golden.json = the output of the reference's driver built with the reference's own headers
    (tests/golden/make_driver_golden.py): directed batches, inserts and deletes, affected-walk counts."""
    gold = driver_golden()
    sizes, counts = replay_throughput_latency(wo, *cora)
    assert sizes == gold["batch_pairs"] == [5, 50, 490]
    assert counts == gold["walks_affected"]


@pytest.mark.gpu
def test_reference_throughput_latency_driver_on_gpu(wharf, wo, cora, tmp_path):
    exe = brc.OUT / "throughput-latency"
    if not exe.exists():
        pytest.skip("reference callers were not built where the snapshot was taken")
    from dynamicgraphrepresentationlearning_b200 import graph_io
    n, off, adj = cora
    graph_io.write_unweighted_graph(tmp_path / "cora.adj", off, adj)
    r = subprocess.run([str(exe), "-f", str(tmp_path / "cora.adj"), "-w", "10", "-l", "80", "-trials", "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Walking model: DEEPWALK" in r.stdout and "Vertices: 2708 Edges: 10556" in r.stdout
    got = [float(x) for x in re.findall(r"Average number of walks affected = ([0-9.e+]+)", r.stdout)]
    batch_sizes = [int(x) for x in re.findall(r"Batch size = \d+ \| (\d+) ", r.stdout)]
    sizes, want = replay_throughput_latency(wo, n, off, adj)
    gold = driver_golden()                                             # the same source built with the reference's own headers
    assert batch_sizes == sizes == gold["batch_pairs"]
    assert got == want == gold["walks_affected"]
    assert "Average time to generate random walks from scratch" in r.stdout


@pytest.mark.gpu
def test_reference_memory_footprint_driver_on_gpu(cora, tmp_path):
    exe = brc.OUT / "memory-footprint"
    if not exe.exists():
        pytest.skip("reference callers were not built where the snapshot was taken")
    from dynamicgraphrepresentationlearning_b200 import graph_io
    n, off, adj = cora
    graph_io.write_unweighted_graph(tmp_path / "cora.adj", off, adj)
    r = subprocess.run([str(exe), "-f", str(tmp_path / "cora.adj"), "-model", "node2vec", "-det", "false"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Walking model: NODE2VEC" in r.stdout and "Vertices: 2708, Edges: 10556" in r.stdout and "Device memory" in r.stdout


@pytest.mark.gpu
def test_reference_own_wharfmh_tests_on_gpu(tmp_path):
    """The reference's own tests/wharfmh.cpp (/root/reference/tests/wharfmh.cpp:56-186 and on: constructors, degrees, destroy,
    edge counts, walk generation, insert/delete batches), compiled UNCHANGED against include/compat + the library, run on the
    device.  Its fixture shells out to ./SNAPtoAdj on data/flickr-graph, a file the reference does not ship: here
    ./SNAPtoAdj is this package's replacement (graph_io, symmetrise/dedup on the device) and the data file is cora's edge
    list from the reference's own experiments/data."""
    import os
    import sys
    import numpy as np
    exe = brc.OUT / "ref_test_wharfmh"
    if not exe.exists():
        pytest.skip("reference callers were not built where the snapshot was taken")
    (tmp_path / "data").mkdir()
    pairs = np.load(ROOT / "tests" / "golden" / "cora_edge_list.npz")["pairs"]
    (tmp_path / "data" / "flickr-graph").write_text("".join(f"{u} {v}\n" for u, v in pairs.tolist()))
    snap = tmp_path / "SNAPtoAdj"
    snap.write_text(f'#!/bin/sh\nPYTHONPATH={ROOT} exec {sys.executable} -m dynamicgraphrepresentationlearning_b200.graph_io "$@"\n')
    snap.chmod(0o755)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=1500, cwd=tmp_path, env=dict(os.environ, MINI_GTEST_FILTER=""))
    lines = [l for l in r.stdout.splitlines() if l.startswith("[")]
    total = [l for l in lines if l.startswith("[  TOTAL   ]")]
    assert total, (r.stdout[-3000:], r.stderr[-3000:])
    ran, failures = (int(x) for x in re.findall(r"(\d+) tests, (\d+) failures", total[-1])[0])
    failed = [l for l in lines if l.startswith("[  FAILED  ]")]
    assert r.returncode == 0 and failures == 0 and not failed, (failed, r.stdout[-3000:], r.stderr[-3000:])
    assert ran >= 10
