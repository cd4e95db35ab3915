"""The ends of the configuration domain: config::walks_per_vertex and config::walk_length are uint8_t in the reference
(config/globals.h:7-8), so 1 and 255 are the extreme shapes.  tests/golden/shape_golden.json holds the UNMODIFIED
reference's corpora for them (tests/golden/make_shape_golden.py): RMAT scale-8, an insert, the delete of the same batch,
another insert.  The oracle is pinned on the CPU, the engine (both affected-walk search modes) on the GPU."""
import hashlib
import json

import numpy as np
import pytest

from conftest import GOLDEN_DIR

GOLD = json.loads((GOLDEN_DIR / "shape_golden.json").read_text())
CASES = [pytest.param(c, id=f"{c['wpv']}x{c['len']}") for c in GOLD["cases"]]


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("case", CASES)
def test_oracle_extreme_shapes(wo, case):
    n = 1 << GOLD["scale"]
    off, adj = wo.rmat_graph(GOLD["scale"], GOLD["edges_half"])
    o = wo.Oracle(n, off, adj, walks_per_vertex=case["wpv"], walk_length=case["len"])
    assert o.generate() == case["phases"][0]["steps"]
    assert sha(o.walks()) == case["phases"][0]["walks_sha256"]
    for (kind, size, seed), ph in zip(GOLD["ops"], case["phases"][1:]):
        b = wo.generate_batch(size, n, seed, False, False)
        assert o.update(b, delete=kind == "del", remove_dups=True) == ph["affected"]
        assert o.m == ph["m"] and sha(o.affected()[0]) == ph["affected_sha256"] and sha(o.walks()) == ph["walks_sha256"]
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("search", ["scan", "index"])
@pytest.mark.parametrize("case", CASES)
def test_engine_extreme_shapes(wharf, wo, case, search):
    n = 1 << GOLD["scale"]
    off, adj = wo.rmat_graph(GOLD["scale"], GOLD["edges_half"])
    cfg = wharf.Config(walks_per_vertex=case["wpv"], walk_length=case["len"], random_walk_model=wharf.DEEPWALK, deterministic_mode=True,
                       search_mode=wharf.SEARCH_SCAN if search == "scan" else wharf.SEARCH_INDEX)
    with wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg) as e:
        assert e.generate_initial_random_walks() == case["phases"][0]["steps"]
        assert e.walks().shape == (n * case["wpv"], case["len"]) and sha(e.walks()) == case["phases"][0]["walks_sha256"]
        assert len(e.walk(n * case["wpv"] - 1).split()) <= case["len"]
        for (kind, size, seed), ph in zip(GOLD["ops"], case["phases"][1:]):
            b = wo.generate_batch(size, n, seed, False, False)
            fn = e.delete_edges_batch if kind == "del" else e.insert_edges_batch
            affected = fn(b.copy(), sorted=False, remove_dups=True, nn=n)
            assert affected.shape[0] == ph["affected"] and e.number_of_edges() == ph["m"]
            assert sha(affected) == ph["affected_sha256"] and sha(e.walks()) == ph["walks_sha256"]


def test_shapes_outside_the_domain_are_refused(wharf, wo):
    off, adj = wo.rmat_graph(8, 1500)
    for bad in (dict(walks_per_vertex=0), dict(walk_length=0), dict(walks_per_vertex=256), dict(walk_length=256)):
        with pytest.raises(wharf.WharfError) as err:                  # checked before any device work: also without a GPU
            wharf.WharfMH(256, int(adj.shape[0]), off, adj, config=wharf.Config(**bad))
        assert err.value.code == wharf._capi.ERR_INVALID


# ---- batch kinds the other goldens do not mix (tests/golden/make_mixed_golden.py) -----------------------------------------
MIXED = json.loads((GOLDEN_DIR / "mixed_golden.json").read_text())["cases"]
MIXED_CASES = [pytest.param(c, id=f"scale{c['scale']}") for c in MIXED]


def mixed_ops(wo, c):
    n = 1 << c["scale"]
    return [(kind, wo.generate_batch(size, n, seed, bool(loops), bool(directed))) for kind, size, seed, loops, directed in c["ops"]]


@pytest.mark.parametrize("case", MIXED_CASES)
def test_oracle_mixed_batches(wo, case):
    """Absent-edge deletes, a directed batch with self-loops, the same batch again, a directed delete and its re-insert."""
    n = 1 << case["scale"]
    off, adj = wo.rmat_graph(case["scale"], case["edges_half"])
    o = wo.Oracle(n, off, adj, walks_per_vertex=case["wpv"], walk_length=case["len"])
    assert o.generate() == case["phases"][0]["steps"] and sha(o.walks()) == case["phases"][0]["walks_sha256"]
    for (kind, b), ph in zip(mixed_ops(wo, case), case["phases"][1:]):
        assert o.update(b, delete=kind == "del", remove_dups=True) == ph["affected"]
        assert o.m == ph["m"] and sha(o.affected()[0]) == ph["affected_sha256"] and sha(o.walks()) == ph["walks_sha256"]
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("search", ["scan", "index", "auto"])
@pytest.mark.parametrize("case", MIXED_CASES)
def test_engine_mixed_batches(wharf, wo, case, search):
    n = 1 << case["scale"]
    off, adj = wo.rmat_graph(case["scale"], case["edges_half"])
    mode = {"scan": wharf.SEARCH_SCAN, "index": wharf.SEARCH_INDEX, "auto": wharf.SEARCH_AUTO}[search]
    cfg = wharf.Config(walks_per_vertex=case["wpv"], walk_length=case["len"], random_walk_model=wharf.NODE2VEC, deterministic_mode=True, search_mode=mode)
    with wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg) as e:
        assert e.generate_initial_random_walks() == case["phases"][0]["steps"] and sha(e.walks()) == case["phases"][0]["walks_sha256"]
        if search == "auto":
            e.build_index()                                       # AUTO uses the index only when one is valid
        for (kind, b), ph in zip(mixed_ops(wo, case), case["phases"][1:]):
            fn = e.delete_edges_batch if kind == "del" else e.insert_edges_batch
            affected = fn(b.copy(), sorted=False, remove_dups=True, nn=n)
            assert affected.shape[0] == ph["affected"] and e.number_of_edges() == ph["m"]
            assert sha(affected) == ph["affected_sha256"] and sha(e.walks()) == ph["walks_sha256"]
