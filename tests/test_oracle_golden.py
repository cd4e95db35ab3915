"""CPU tests: the plain-C oracle against the golden vectors taken from the unmodified reference
(tests/golden/make_golden.py) and the KATs quoted in SURVEY.md section 8c."""
import hashlib

import numpy as np
import pytest

from conftest import GOLDEN_DIR, SIX_EDGES, SIX_OFFSETS


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_random_kats(wo, golden):
    # utility::Random (utils/utility.h:152-223); SURVEY 8c quotes the first three draws of seeds 0, 1, 2
    assert wo.random_stream(0, 3) == [10407335079877134008, 3962074050977524353, 12330921719341810270]
    assert wo.random_stream(1, 3) == [7437716154062571297, 1116485288419732103, 9741853272534875990]
    assert wo.random_stream(2, 3) == [2528678525309737254, 8345201964650685463, 16366336586872171348]
    for seed, vals in golden["random"].items():
        assert wo.random_stream(int(seed), len(vals)) == vals
    for seed, vals in golden["drand"].items():
        assert wo.drand_stream(int(seed), len(vals)) == vals


def test_pairing_kats(wo, golden):
    L = wo.lib()
    import ctypes as C
    assert L.wo_pair32(65535, 65535) == 4294967295          # tests/pairings.cpp:29-37
    assert L.wo_pair_triplet32(123, 25, 200) == 229643916   # tests/pairings.cpp:75-90
    for x, y, z in golden["pair32"]:
        assert L.wo_pair32(x, y) == z
        a, b = C.c_uint32(), C.c_uint32()
        L.wo_unpair32(z, C.byref(a), C.byref(b))
        assert (a.value, b.value) == (x, y)
    for x, y, z in golden["pair64"]:
        assert L.wo_pair64(x, y) == z
    for z, x, y in golden["unpair64"]:
        a, b = C.c_uint64(), C.c_uint64()
        L.wo_unpair64(z, C.byref(a), C.byref(b))
        assert (a.value, b.value) == (x, y)
    for a_, b_, c_, z in golden["triplet64"]:
        assert L.wo_pair_triplet64(a_, b_, c_) == z
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        L.wo_unpair_triplet64(z, C.byref(a), C.byref(b), C.byref(c))
        assert (a.value, b.value, c.value) == (a_, b_, c_)


def test_pairing_round_trips(wo):
    # tests/pairings.cpp:44-73 and :95-126 domains (sample of 20000 instead of 10 M)
    import ctypes as C
    L = wo.lib()
    rng = np.random.default_rng(7)
    for x, y in rng.integers(0, 65536, (20000, 2)):
        a, b = C.c_uint32(), C.c_uint32()
        L.wo_unpair32(L.wo_pair32(int(x), int(y)), C.byref(a), C.byref(b))
        assert (a.value, b.value) == (x, y)
    for x, y, z in zip(rng.integers(0, 65536, 20000), rng.integers(0, 65536, 20000), rng.integers(0, 2**32 - 1, 20000)):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        L.wo_unpair_triplet64(L.wo_pair_triplet64(int(x), int(y), int(z)), C.byref(a), C.byref(b), C.byref(c))
        assert (a.value, b.value, c.value) == (x, y, z)


def test_batch_generator(wo, golden):
    for g in golden["genbatch"]:
        b = wo.generate_batch(g["edges"], g["vertices"], g["seed"], bool(g["self_loops"]), bool(g["directed"]))
        assert b.shape[0] == g["count"]
        assert b[0].tolist() == g["first"] and b[-1].tolist() == g["last"]
        assert sha(b) == g["sha256"]


def test_six_vertex_corpus(wo, golden):
    six = golden["six"]
    o = wo.Oracle(6, SIX_OFFSETS, SIX_EDGES, walks_per_vertex=2, walk_length=8)
    o.generate()
    # SURVEY 8c quotes these rows verbatim
    assert o.walks()[0].tolist() == [0, 1, 0, 1, 2, 1, 2, 0]
    assert o.walks()[11].tolist() == [5, 2, 4, 2, 5, 2, 5, 2]
    assert o.walks().tolist() == six["walks"][0]
    batch = np.array(six["batch"], dtype=np.uint32)
    assert o.update(batch, delete=False, remove_dups=True) == 12
    assert o.walks().tolist() == six["walks"][1]
    assert o.walks()[2].tolist() == [2, 4, 5, 2, 0, 2, 1, 2]
    assert o.update(batch, delete=True, remove_dups=True) == 12
    assert o.walks().tolist() == six["walks"][2]
    assert o.walks()[2].tolist() == [2, 4, 2, 4, 2, 3, 2, 0]


def test_cora(wo, golden, cora):
    n, off, adj = cora
    g = golden["cora"]
    o = wo.Oracle(n, off, adj)
    assert o.generate() == g["phases"][0]["steps"] == 2166400
    w = o.walks()
    assert w[0, :10].tolist() == g["walk0_prefix"] == [0, 435, 258, 268, 616, 438, 710, 847, 1032, 848]
    assert w[2708, :10].tolist() == g["walk2708_prefix"] == [0, 258, 268, 232, 14, 435, 751, 258, 8, 14]
    assert sha(w) == g["phases"][0]["walks_sha256"]
    assert wo.corpus_hash(w) == g["phases"][0]["corpus_fnv1a"]
    b = wo.generate_batch(100, n, 0, False, False)
    assert b.shape[0] == 198 and b[0].tolist() == [0, 836] and b[-1].tolist() == [2005, 852]
    for ph, delete in zip(g["phases"][1:], (False, True)):
        assert o.update(b, delete=delete, remove_dups=False) == ph["affected"]
        assert o.m == ph["m"]
        assert sha(o.affected()[0]) == ph["affected_sha256"]
        assert sha(o.walks()) == ph["walks_sha256"]
    assert g["phases"][1]["affected"] == 23906 and g["phases"][1]["m"] == 10752


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_rmat_sequences(wo, golden, idx):
    g = golden["rmat"][idx]
    n = 1 << g["scale"]
    off, adj = wo.rmat_graph(g["scale"], g["edges_half"], g["graph_seed"])
    model = wo.NODE2VEC if g["model"] == "node2vec" else wo.DEEPWALK
    o = wo.Oracle(n, off, adj, walks_per_vertex=g["wpv"], walk_length=g["len"], model=model)
    assert o.generate() == g["phases"][0]["steps"]
    assert o.m == g["phases"][0]["m"]
    assert sha(o.walks()) == g["phases"][0]["walks_sha256"]
    for (kind, bs, seed), ph in zip(g["ops"], g["phases"][1:]):
        b = wo.generate_batch(bs, n, seed, False, False)
        assert o.update(b, delete=(kind == "del"), remove_dups=True) == ph["affected"]
        assert o.m == ph["m"]
        assert sha(o.affected()[0]) == ph["affected_sha256"]
        assert sha(o.walks()) == ph["walks_sha256"]
    if g["scale"] == 8:
        z = np.load(GOLDEN_DIR / "rmat8_walks.npz")
        assert np.array_equal(z[f"walks{len(g['ops'])}"], o.walks())


def test_deterministic_models_agree(wo):
    # in deterministic mode DeepWalk and node2vec produce identical corpora
    off, adj = wo.rmat_graph(9, 3000)
    a = wo.Oracle(512, off, adj, model=wo.DEEPWALK, walk_length=20, walks_per_vertex=3)
    b = wo.Oracle(512, off, adj, model=wo.NODE2VEC, walk_length=20, walks_per_vertex=3)
    a.generate(), b.generate()
    assert np.array_equal(a.walks(), b.walks())


def test_compressed_find_next(wo):
    import ctypes as C
    L = wo.lib()
    codes = np.array([L.wo_pair64(w * 8 + p, nxt) for w, p, nxt in [(0, 1, 5), (3, 7, 0xFFFFFFFE), (2, 0, 9)]], dtype=np.uint64)
    f = lambda w, p: L.wo_compressed_find_next(C.c_void_p(codes.ctypes.data), 3, w, p, 8)
    assert f(0, 1) == 5 and f(3, 7) == 0xFFFFFFFE and f(2, 0) == 9 and f(1, 1) == 0xFFFFFFFF
