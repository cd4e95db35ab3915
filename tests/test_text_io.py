"""Ingest and export either side of the path (dynamicgraphrepresentationlearning_b200/graph_io.py,
wmh_parse_* / wmh_format_* / wmh_write_walks in include/wharfmh_b200.h).

Pinned to the reference's own tools: tests/golden/io_golden.json holds what experiments/bin/SNAPtoAdj -s writes for
experiments/data/cora-graph (tests/golden/make_io_golden.py), tests/golden/golden.json the FNV-1a of the text the
unmodified reference prints for the cora corpus (WharfMH::walk of every walk + newline).  The text functions are host
code and run without a GPU; the set work of ingest is the engine's batch insert and is marked gpu."""
import hashlib
import json

import numpy as np
import pytest

from conftest import GOLDEN_DIR


@pytest.fixture(scope="module")
def gio(wharf):
    from dynamicgraphrepresentationlearning_b200 import graph_io
    return graph_io


@pytest.fixture(scope="module")
def io_golden():
    return json.loads((GOLDEN_DIR / "io_golden.json").read_text())


@pytest.fixture(scope="module")
def cora_pairs():
    return np.load(GOLDEN_DIR / "cora_edge_list.npz")["pairs"]


def fnv1a(data: bytes) -> int:
    h = 0xCBF29CE484222325
    for c in data:
        h = ((h ^ c) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h


def edge_list_text(pairs) -> bytes:
    return "".join(f"{u} {v}\n" for u, v in pairs.tolist()).encode()


# ---- host text functions (no GPU) -----------------------------------------------------------------------------------
def test_adjacency_graph_text_equals_snaptoadj(gio, io_golden, cora, tmp_path):
    n, off, adj = cora
    text = gio.format_unweighted_graph(off, adj)
    assert len(text) == io_golden["cora_adj_bytes"] and hashlib.sha256(text).hexdigest() == io_golden["cora_adj_sha256"]
    n2, m2, off2, adj2 = gio.read_unweighted_graph(text)
    assert (n2, m2) == (n, adj.shape[0]) and off2.dtype == np.uint64 and adj2.dtype == np.uint32
    assert np.array_equal(off2, off) and np.array_equal(adj2, adj)          # n offsets, no n+1-th (IO.h:97)
    path = tmp_path / "cora.adj"
    gio.write_unweighted_graph(path, np.append(off, np.uint64(adj.shape[0])), adj, row_ptr=True)
    assert path.read_bytes() == text
    assert gio.read_unweighted_graph(path)[0] == n
    toy = gio.read_unweighted_graph(io_golden["toy_adj"].encode())
    assert toy[:2] == (5, 6) and toy[2].tolist() == [0, 2, 3, 3, 5] and toy[3].tolist() == [3, 4, 3, 0, 1, 0]
    assert gio.format_unweighted_graph(toy[2], toy[3]).decode() == io_golden["toy_adj"]
    assert gio.read_unweighted_graph(b"  AdjacencyGraph\t0\r\n0\n")[:2] == (0, 0)
    assert gio.format_unweighted_graph(np.empty(0), np.empty(0)) == b"AdjacencyGraph\n0\n0\n"


@pytest.mark.parametrize("text,what", [
    (b"Adjacency 1 0 0", "not an AdjacencyGraph"), (b"AdjacencyGraphX 0 0", "not an AdjacencyGraph"), (b"", "not an AdjacencyGraph"),
    (b"AdjacencyGraph", "vertex count"), (b"AdjacencyGraph 3", "edge count"),
    (b"AdjacencyGraph 2 1 0 1", "fewer than n"), (b"AdjacencyGraph 2 1 0 1 1 1", "more than n"),
    (b"AdjacencyGraph 2 1 1 0 1", "non-decreasing"), (b"AdjacencyGraph 2 1 0 2 1", "non-decreasing"),
    (b"AdjacencyGraph 2 1 0 1 2", "neighbour id >= n"), (b"AdjacencyGraph 2 1 0 x 1", "fewer than")])
def test_adjacency_graph_rejects_bad_files(wharf, gio, text, what):
    with pytest.raises(wharf.WharfError, match=what):              # reference: assert / garbage (IO.h:87,94)
        gio.read_unweighted_graph(text)


def test_edge_list_parser(wharf, gio, io_golden, cora_pairs, tmp_path):
    text = edge_list_text(cora_pairs)
    assert hashlib.sha256(text).hexdigest() == io_golden["cora_edge_list_sha256"]       # the fixture IS the reference's file
    pairs, vertices = gio.read_edge_list(text)
    assert vertices == 2708 and pairs.dtype == np.uint32 and np.array_equal(pairs, cora_pairs)
    path = tmp_path / "cora-graph"
    path.write_bytes(text)
    assert np.array_equal(gio.read_edge_list(path)[0], cora_pairs)
    pairs, vertices = gio.read_edge_list(b"# SNAP header\n3 1\r\n\n1\t3\n  2 2  \n%x\n0 3")     # no final newline
    assert pairs.tolist() == [[3, 1], [1, 3], [2, 2], [0, 3]] and vertices == 4
    assert gio.read_edge_list(io_golden["toy_edge_list"].encode())[0].tolist() == [[3, 1], [1, 3], [2, 2], [0, 3], [0, 3], [4, 0]]
    empty, vertices = gio.read_edge_list(b"# nothing\n\n")
    assert empty.shape == (0, 2) and vertices == 0
    assert gio.read_edge_list(b"4294967293 0\n")[1] == 4294967294
    for bad, code in ((b"1 2 3\n", 1), (b"1\n", 1), (b"a b\n", 1), (b"1 -2\n", 1), (b"0 1\n1 4294967295\n", 5), (b"1 99999999999999999999999\n", 1)):
        with pytest.raises(wharf.WharfError) as err:
            gio.read_edge_list(bad)
        assert err.value.code == code
    with pytest.raises(wharf.WharfError, match="line 3"):
        gio.read_edge_list(b"# c\n0 1\n0 x\n")
    # the C entry point reports the count when the buffer is too small
    import ctypes as C
    lib = wharf._capi.load()
    k, v = C.c_uint64(), C.c_uint64()
    small = np.zeros((2, 2), dtype=np.uint32)
    rc = lib.wmh_parse_edge_list(b"0 1\n2 3\n4 5\n", 12, small.ctypes.data, 2, C.byref(k), C.byref(v))
    assert rc == wharf._capi.ERR_RANGE and k.value == 3 and small.tolist() == [[0, 1], [2, 3]]


def test_walk_text_equals_reference_corpus_text(gio, golden, cora, wo):
    n, off, adj = cora
    o = wo.Oracle(n, off, adj, walks_per_vertex=10, walk_length=80)
    o.generate()
    text = gio.format_walks(o.walks())
    assert fnv1a(text) == golden["cora"]["phases"][0]["corpus_fnv1a"]       # what the unmodified reference prints
    assert text.startswith(b"0 435 258 268 616 438 710 847 1032 848 ") and text.count(b"\n") == 10 * n
    o.close()
    rows = np.array([[1, 2, 0xFFFFFFFF], [0xFFFFFFFF] * 3, [4294967294, 0, 7]], dtype=np.uint32)
    assert gio.format_walks(rows) == b"1 2 \n\n4294967294 0 7 \n"             # a stopped walker ends its line early
    assert gio.format_walks(np.empty((0, 80), dtype=np.uint32)) == b""
    rng = np.random.default_rng(3)
    rows = rng.integers(0, 2 ** 32 - 1, size=(200, 17), dtype=np.uint64).astype(np.uint32)
    assert gio.format_walks(rows).decode() == "".join("".join(f"{v} " for v in r) + "\n" for r in rows.tolist())


def test_create_edge_stream_follows_the_reference_loop(gio, cora_pairs, tmp_path):
    def reference_loop(pairs, eps):                # vertex-classification.cpp:14-27, restated
        temp = []
        for it, (a, b) in enumerate(pairs.tolist()):
            if it % eps == 0:
                temp.append([])
            temp[-1].append((a, b))
            temp[-1].append((b, a))
        return temp
    for eps in (5000, 1000, 5429, 7):
        stream = gio.create_edge_stream(cora_pairs, eps)
        want = reference_loop(cora_pairs, eps)
        assert len(stream) == len(want)
        assert all(s.dtype == np.uint32 and s.tolist() == [list(p) for p in w] for s, w in zip(stream, want))
    path = tmp_path / "stream"
    path.write_bytes(edge_list_text(cora_pairs[:11]))
    assert [s.shape[0] for s in gio.create_edge_stream(path, 4)] == [8, 8, 6]
    assert gio.create_edge_stream(np.empty((0, 2), dtype=np.uint32)) == []
    with pytest.raises(ValueError):
        gio.create_edge_stream(cora_pairs, 0)


# ---- through the engine (GPU) ---------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_snap_to_adj_equals_the_reference_tool(wharf, gio, io_golden, cora, cora_pairs):
    n, off, adj = cora
    got_n, got_m, got_off, got_adj = gio.snap_to_adj(cora_pairs)
    assert (got_n, got_m) == (n, adj.shape[0]) and np.array_equal(got_off, off) and np.array_equal(got_adj, adj)
    assert hashlib.sha256(gio.format_unweighted_graph(got_off, got_adj)).hexdigest() == io_golden["cora_adj_sha256"]
    _, _, toy_off, toy_adj = gio.snap_to_adj(io_golden["toy_edge_list"].encode())            # comment, loop, duplicates
    assert gio.format_unweighted_graph(toy_off, toy_adj).decode() == io_golden["toy_adj"]
    # directed: rows are sets here (WharfMH needs unique rows, wharfmh.h:58); loops dropped
    _, m, d_off, d_adj = gio.snap_to_adj(cora_pairs, symmetric=False)
    keys = np.unique(cora_pairs[:, 0].astype(np.int64) << 32 | cora_pairs[:, 1].astype(np.int64))
    assert m == keys.shape[0] and np.array_equal(np.repeat(np.arange(n), np.diff(np.append(d_off, np.uint64(m)).astype(np.int64))) << 32 | d_adj.astype(np.int64), keys)
    assert gio.snap_to_adj(np.empty((0, 2), dtype=np.uint32))[:2] == (0, 0)


@pytest.mark.gpu
def test_stream_ingest_and_walks_txt(wharf, gio, golden, cora, cora_pairs, wo, tmp_path):
    """The vertex-classification pipeline's shape (vertex-classification.cpp:136-197): a graph grown from an edge stream,
    the walks each batch changed handed over as text."""
    n, off, adj = cora
    cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.DEEPWALK, deterministic_mode=True)
    # the whole file as one stream == SNAPtoAdj -s; its corpus text == what the unmodified reference prints
    with gio.graph_from_edge_list(cora_pairs, config=cfg) as e:
        assert e.number_of_vertices() == n and e.number_of_edges() == adj.shape[0]
        e.generate_initial_random_walks()
        path = tmp_path / "walks.txt"
        written = gio.write_walks(e, path, np.arange(10 * n, dtype=np.uint64))
        text = path.read_bytes()
        assert written == len(text) and fnv1a(text) == golden["cora"]["phases"][0]["corpus_fnv1a"]
        b = wo.generate_batch(100, n, 0, False, False)
        affected = e.insert_edges_batch(b.copy(), sorted=True, remove_dups=False)
        assert gio.write_walks(e, path, affected) > 0
        lines = path.read_bytes().split(b"\n")
        assert len(lines) == affected.shape[0] + 1 and lines[-1] == b""
        for k in (0, 1, affected.shape[0] // 2, affected.shape[0] - 1):
            assert lines[k].decode() == e.walk(int(affected[k]))                      # WharfMH::walk + endl per affected walk
        before = path.stat().st_size
        gio.write_walks(e, path, affected[:3], append=True)
        assert path.read_bytes()[before:] == b"".join(lines[:3][i] + b"\n" for i in range(3))
        assert fnv1a(gio.format_walks(e.walks())) == golden["cora"]["phases"][1]["corpus_fnv1a"]
        with pytest.raises(wharf.WharfError):
            gio.write_walks(e, path, np.array([10 * n], dtype=np.uint64))             # not a walk of this engine
    # grown batch by batch, the graph ends the same; a sharded engine writes only walks it holds
    with wharf.WharfMH(n, config=cfg) as e:
        with pytest.raises(wharf.WharfError):
            gio.write_walks(e, tmp_path / "none.txt", np.array([0], dtype=np.uint64))  # no corpus yet
        for batch in gio.create_edge_stream(cora_pairs, 1000):
            e.insert_edges_batch(batch, sorted=False, remove_dups=True)
        rp, col = e.flatten_graph()
        assert np.array_equal(rp[:-1], off) and np.array_equal(col, adj)
    shard = wharf.Config(walks_per_vertex=10, walk_length=80, deterministic_mode=True, rank=1, world=2)
    with wharf.WharfMH(n, adj.shape[0], off, adj, config=shard) as e:
        e.generate_initial_random_walks()
        mine = e.local_walk_ids()
        assert gio.write_walks(e, tmp_path / "shard.txt", mine[:5]) > 0
        with pytest.raises(wharf.WharfError, match="not held"):
            gio.write_walks(e, tmp_path / "shard.txt", np.array([0], dtype=np.uint64))


# ---- a second real graph: experiments/data/wiki-graph (1996 of its 17981 lines are self-loops, 42 vertices end up isolated) ----
@pytest.fixture(scope="module")
def wiki():
    gold = json.loads((GOLDEN_DIR / "wiki_golden.json").read_text())
    z = np.load(GOLDEN_DIR / "wiki_graph.npz")
    return gold, np.load(GOLDEN_DIR / "wiki_edge_list.npz")["pairs"], z["offsets"], z["edges"]


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_wiki_text_and_oracle_follow_the_reference(gio, wo, wiki):
    gold, pairs, off, adj = wiki
    n = gold["n"]
    assert (n, adj.shape[0], pairs.shape[0], int((pairs[:, 0] == pairs[:, 1]).sum())) == (2405, gold["m"], gold["lines"], gold["self_loops"])
    assert hashlib.sha256(gio.format_unweighted_graph(off, adj)).hexdigest() == gold["adj_sha256"]          # SNAPtoAdj -s's file
    got, vertices = gio.read_edge_list("".join(f"{u}\t{v}\n" for u, v in pairs.tolist()).encode())         # the file is tab-separated
    assert vertices == n and np.array_equal(got, pairs)
    # the CPU restatement against the unmodified reference's corpora on this graph (isolated starts included)
    o = wo.Oracle(n, off, adj)
    assert o.generate() == gold["phases"][0]["steps"] == 80 * 10 * int((np.diff(np.append(off, np.uint64(adj.shape[0])).astype(np.int64)) > 0).sum())
    assert o.walks()[0, :10].tolist() == gold["walk0_prefix"]
    assert sha(o.walks()) == gold["phases"][0]["walks_sha256"]
    for (kind, size, seed), ph in zip(gold["ops"], gold["phases"][1:]):
        b = wo.generate_batch(size, n, seed, False, False)
        assert o.update(b, delete=kind == "del", remove_dups=True) == ph["affected"]
        assert o.m == ph["m"] and sha(o.affected()[0]) == ph["affected_sha256"] and sha(o.walks()) == ph["walks_sha256"]
    o.close()


@pytest.mark.gpu
def test_wiki_ingest_and_corpora_on_gpu(wharf, gio, wo, wiki):
    gold, pairs, off, adj = wiki
    n = gold["n"]
    got_n, got_m, got_off, got_adj = gio.snap_to_adj(pairs)                         # loops and duplicates dropped on the device
    assert (got_n, got_m) == (n, gold["m"]) and np.array_equal(got_off, off) and np.array_equal(got_adj, adj)
    cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.DEEPWALK, deterministic_mode=True)
    for mode in (wharf.SEARCH_SCAN, wharf.SEARCH_INDEX):
        cfg.search_mode = mode
        with gio.graph_from_edge_list(pairs, config=cfg) as e:                      # straight from the edge list
            assert e.generate_initial_random_walks() == gold["phases"][0]["steps"]
            assert sha(e.walks()) == gold["phases"][0]["walks_sha256"]
            assert fnv1a(gio.format_walks(e.walks()[:500])) == fnv1a("".join(e.walk(w) + "\n" for w in range(500)).encode())
            for (kind, size, seed), ph in zip(gold["ops"], gold["phases"][1:]):
                b = wo.generate_batch(size, n, seed, False, False)
                fn = e.delete_edges_batch if kind == "del" else e.insert_edges_batch
                affected = fn(b.copy(), sorted=False, remove_dups=True, nn=4096)
                assert affected.shape[0] == ph["affected"] and e.number_of_edges() == ph["m"]
                assert sha(affected) == ph["affected_sha256"] and sha(e.walks()) == ph["walks_sha256"]


@pytest.mark.gpu
def test_generate_batch_of_edges_equals_the_reference_generator(wharf, golden):
    """utility::generate_batch_of_edges through the host-facing entry point (what throughput-latency.cpp:121 calls):
    byte-identical to the unmodified reference's batches (tests/golden/golden.json "genbatch")."""
    for g in golden["genbatch"]:
        b = wharf.generate_batch_of_edges(g["edges"], g["vertices"], g["seed"], bool(g["self_loops"]), bool(g["directed"]))
        assert b.dtype == np.uint32 and b.shape == (g["count"], 2)
        assert b[0].tolist() == g["first"] and b[-1].tolist() == g["last"]
        assert hashlib.sha256(b.tobytes()).hexdigest() == g["sha256"]
    with pytest.raises(wharf.WharfError):
        wharf.generate_batch_of_edges(10, 1, 0)                       # fewer than two vertices


# ---- the incremental flow of vertex-classification.cpp:118-197: vertices in isolation, then the edge stream ----------
@pytest.fixture(scope="module")
def stream_golden():
    return json.loads((GOLDEN_DIR / "stream_golden.json").read_text())


def test_edge_stream_from_empty_graph_oracle(gio, wo, stream_golden, cora_pairs):
    g = stream_golden
    n = g["n"]
    o = wo.Oracle(n, np.zeros(n, dtype=np.uint64), np.zeros(0, dtype=np.uint32))
    assert o.generate() == g["phases"][0]["steps"] == 0                # every walk is its start vertex alone
    assert sha(o.walks()) == g["phases"][0]["walks_sha256"]
    stream = gio.create_edge_stream(cora_pairs, g["partition"])
    assert len(stream) == len(g["phases"]) - 1
    for batch, ph in zip(stream, g["phases"][1:]):
        assert o.update(batch.copy(), delete=False, remove_dups=True) == ph["affected"]
        assert o.m == ph["m"] and sha(o.affected()[0]) == ph["affected_sha256"] and sha(o.walks()) == ph["walks_sha256"]
    o.close()


@pytest.mark.gpu
def test_edge_stream_from_empty_graph_on_gpu(wharf, gio, stream_golden, cora, cora_pairs, tmp_path):
    g = stream_golden
    n, off, adj = cora
    cfg = wharf.Config(walks_per_vertex=10, walk_length=80, random_walk_model=wharf.DEEPWALK, deterministic_mode=True)
    with wharf.WharfMH(n, config=cfg) as e:                              # WharfMH(n, m): all vertices in isolation (wharfmh.h:26)
        assert e.generate_initial_random_walks() == 0
        assert sha(e.walks()) == g["phases"][0]["walks_sha256"]
        assert gio.format_walks(e.walks()[:3]) == b"0 \n1 \n2 \n"
        for batch, ph in zip(gio.create_edge_stream(cora_pairs, g["partition"]), g["phases"][1:]):
            affected = e.insert_edges_batch(batch, sorted=False, remove_dups=True)     # vertex-classification.cpp:170
            assert affected.shape[0] == ph["affected"] and e.number_of_edges() == ph["m"]
            assert sha(affected) == ph["affected_sha256"] and sha(e.walks()) == ph["walks_sha256"]
            e.write_walks(tmp_path / "walks.txt", affected)                            # :172-178
            assert fnv1a((tmp_path / "walks.txt").read_bytes()[:4096]) == fnv1a(gio.format_walks(e.walks_gather(affected[:200].astype(np.uint32)))[:4096])
        rp, col = e.flatten_graph()
        assert np.array_equal(rp[:-1], off) and np.array_equal(col, adj)                # the stream rebuilt SNAPtoAdj -s's graph


# ---- differential fuzz of the text parsers against a line-by-line Python restatement ------------------------------------
def py_edge_list(text: bytes):
    """What wmh_parse_edge_list accepts: per line (split at \\n) - blank or first non-blank char # / % -> skipped; else exactly two
    unsigned decimals < 2^32 - 2 separated by blanks.  Returns the pairs or None when any line is malformed."""
    pairs = []
    for line in text.split(b"\n"):
        body = line.strip(b" \t\r\f\v")
        if not body or body[:1] in (b"#", b"%"):
            continue
        tok = body.split()
        if len(tok) != 2 or not all(t.isdigit() for t in tok):
            return None
        a, b = int(tok[0]), int(tok[1])
        if a >= 2 ** 64 or b >= 2 ** 64 or a >= 0xFFFFFFFE or b >= 0xFFFFFFFE:
            return None
        pairs.append([a, b])
    return pairs


def test_edge_list_parser_differential_fuzz(wharf, gio):
    rng = np.random.default_rng(2024)
    alphabet = np.frombuffer(b"0123456789  \t\n\n\r#%x-", dtype=np.uint8)
    agree = {"ok": 0, "bad": 0}
    for trial in range(4000):
        text = rng.choice(alphabet, size=int(rng.integers(0, 40))).tobytes()
        if trial % 3 == 0:                                              # mostly well-formed files with one flaw somewhere
            lines = [f"{rng.integers(0, 10 ** int(rng.integers(1, 11)))}{' ' * int(rng.integers(1, 3))}{rng.integers(0, 5000)}" for _ in range(int(rng.integers(0, 6)))]
            text = ("\n".join(lines) + ("\n" if rng.random() < 0.5 else "")).encode() + (text if rng.random() < 0.3 else b"")
        want = py_edge_list(text)
        try:
            got, vertices = gio.read_edge_list(text)
        except wharf.WharfError:
            assert want is None, text
            agree["bad"] += 1
            continue
        assert want is not None and got.tolist() == want, text
        assert vertices == (max(max(p) for p in want) + 1 if want else 0)
        agree["ok"] += 1
    assert agree["ok"] > 300 and agree["bad"] > 300


def test_adjacency_parser_fuzz_never_reads_past_the_text(wharf, gio):
    rng = np.random.default_rng(7)
    base = b"AdjacencyGraph\n3\n4\n0\n2\n3\n1\n2\n0\n0\n"
    assert gio.read_unweighted_graph(base)[:2] == (3, 4)
    for _ in range(3000):
        cut = base[: int(rng.integers(0, len(base) + 1))]
        noise = bytes(rng.choice(np.frombuffer(b"0123456789 \n9x", dtype=np.uint8), size=int(rng.integers(0, 6))))
        text = cut + noise if rng.random() < 0.5 else cut[: len(cut) // 2] + noise + cut[len(cut) // 2:]
        try:
            n, m, off, adj = gio.read_unweighted_graph(text)
        except (wharf.WharfError, MemoryError, ValueError):
            continue
        tok = text.split()
        assert tok[0] == b"AdjacencyGraph" and len(tok) == n + m + 3 and off.tolist() == [int(t) for t in tok[3:3 + n]] and adj.tolist() == [int(t) for t in tok[3 + n:]]
