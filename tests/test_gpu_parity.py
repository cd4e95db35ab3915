"""GPU parity tests: the CUDA path (through the C ABI) against the golden vectors of the unmodified reference
and against the CPU oracle on the same seeded inputs.  Integer work: the bar is bit-exact."""
import hashlib

import numpy as np
import pytest

from conftest import GOLDEN_DIR, SIX_EDGES, SIX_OFFSETS

pytestmark = pytest.mark.gpu


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def make(wharf, n, off, adj, **kw):
    cfg = wharf.Config(**kw)
    return wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg)


def test_six_vertex_golden(wharf, golden):
    six = golden["six"]
    with make(wharf, 6, SIX_OFFSETS, SIX_EDGES, walks_per_vertex=2, walk_length=8, random_walk_model=wharf.DEEPWALK) as e:
        assert e.generate_initial_random_walks() == 96
        assert e.walks().tolist() == six["walks"][0]
        assert e.walk(0) == "0 1 0 1 2 1 2 0 "
        batch = np.array(six["batch"], dtype=np.uint32)
        affected = e.insert_edges_batch(batch.copy(), sorted=False, remove_dups=True, nn=8)
        assert affected.tolist() == list(range(12))
        assert e.number_of_edges() == 20
        assert e.walks().tolist() == six["walks"][1]
        affected = e.delete_edges_batch(batch.copy(), sorted=False, remove_dups=True, nn=8)
        assert affected.shape[0] == 12 and e.number_of_edges() == 18
        assert e.walks().tolist() == six["walks"][2]
        assert e.walk(2) == "2 4 2 4 2 3 2 0 "


@pytest.mark.parametrize("search", ["scan", "index"])
def test_cora_golden(wharf, golden, cora, wo, search):
    n, off, adj = cora
    g = golden["cora"]
    mode = wharf.SEARCH_SCAN if search == "scan" else wharf.SEARCH_INDEX
    with make(wharf, n, off, adj, random_walk_model=wharf.DEEPWALK, search_mode=mode) as e:
        assert e.generate_initial_random_walks() == g["phases"][0]["steps"]
        w = e.walks()
        assert sha(w) == g["phases"][0]["walks_sha256"]
        assert wo.corpus_hash(w) == g["phases"][0]["corpus_fnv1a"]
        b = wo.generate_batch(100, n, 0, False, False)
        for ph, delete in zip(g["phases"][1:], (False, True)):
            fn = e.delete_edges_batch if delete else e.insert_edges_batch
            affected = fn(b.copy(), sorted=True, remove_dups=False)
            assert affected.shape[0] == ph["affected"]
            assert e.number_of_edges() == ph["m"]
            assert sha(affected) == ph["affected_sha256"]
            assert sha(e.walks()) == ph["walks_sha256"]


@pytest.mark.parametrize("idx,search", [(0, "scan"), (0, "index"), (1, "index"), (2, "scan"), (3, "scan"), (3, "index")])
def test_rmat_golden_sequences(wharf, golden, wo, idx, search):
    g = golden["rmat"][idx]
    n = 1 << g["scale"]
    off, adj = wo.rmat_graph(g["scale"], g["edges_half"], g["graph_seed"])
    model = wharf.NODE2VEC if g["model"] == "node2vec" else wharf.DEEPWALK
    mode = wharf.SEARCH_SCAN if search == "scan" else wharf.SEARCH_INDEX
    with make(wharf, n, off, adj, walks_per_vertex=g["wpv"], walk_length=g["len"], random_walk_model=model, search_mode=mode) as e:
        assert e.generate_initial_random_walks() == g["phases"][0]["steps"]
        assert sha(e.walks()) == g["phases"][0]["walks_sha256"]
        for (kind, bs, seed), ph in zip(g["ops"], g["phases"][1:]):
            b = wo.generate_batch(bs, n, seed, False, False)
            fn = e.delete_edges_batch if kind == "del" else e.insert_edges_batch
            affected = fn(b.copy(), sorted=False, remove_dups=True, nn=n)
            assert affected.shape[0] == ph["affected"]
            assert e.number_of_edges() == ph["m"]
            assert sha(affected) == ph["affected_sha256"]
            assert sha(e.walks()) == ph["walks_sha256"]
        if g["scale"] == 8:
            z = np.load(GOLDEN_DIR / "rmat8_walks.npz")
            assert np.array_equal(z[f"walks{len(g['ops'])}"], e.walks())


def random_batch(rng, n, k, symmetric=True):
    b = rng.integers(0, n, (k, 2)).astype(np.uint32)
    if symmetric:
        b = np.concatenate([b, b[:, ::-1]])
    return np.ascontiguousarray(b)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_update_sequences_vs_oracle(wharf, wo, seed):
    """Unsorted batches with duplicates, self-loops, already-present and absent edges; CSR, affected set,
    earliest positions and corpus must all equal the oracle's after every batch."""
    rng = np.random.default_rng(seed)
    scale = 9 + seed
    n = 1 << scale
    off, adj = wo.rmat_graph(scale, 6 * n, seed=seed + 3)
    wpv, L = 4 + seed, 17 + 5 * seed
    o = wo.Oracle(n, off, adj, walks_per_vertex=wpv, walk_length=L)
    mode = [wharf.SEARCH_SCAN, wharf.SEARCH_INDEX, wharf.SEARCH_AUTO][seed]
    with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=L, search_mode=mode) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert np.array_equal(e.walks(), o.walks())
        if mode == wharf.SEARCH_AUTO:
            e.build_index()
        for step in range(6):
            delete = step % 3 == 2
            remove_dups = step % 2 == 0
            if delete:
                row_ptr, col = o.csr()
                pick = rng.integers(0, col.shape[0], 40)
                src = np.searchsorted(row_ptr, pick, side="right") - 1
                b = np.stack([src, col[pick]], 1).astype(np.uint32)
                b = np.concatenate([b, b[:, ::-1], random_batch(rng, n, 5)])
            else:
                b = random_batch(rng, n, 3 + 20 * step)
                b = np.concatenate([b, b[:7], np.array([[5, 5], [9, 9]], dtype=np.uint32)])
            rng.shuffle(b)
            expect_sorted = b[np.lexsort((b[:, 1], b[:, 0]))]
            mine = np.ascontiguousarray(b.copy())
            na = o.update(b, delete=delete, remove_dups=remove_dups)
            fn = e.delete_edges_batch if delete else e.insert_edges_batch
            affected = fn(mine, sorted=False, remove_dups=remove_dups)
            assert np.array_equal(mine, expect_sorted)            # caller's array sorted in place (wharfmh.h:452)
            ids, pos = e.affected_walks()
            assert affected.shape[0] == na
            assert np.array_equal(ids, o.affected()[0]) and np.array_equal(pos, o.affected()[1])
            assert e.number_of_edges() == o.m
            rp, cl = e.flatten_graph()
            orp, ocl = o.csr()
            assert np.array_equal(rp, orp) and np.array_equal(cl, ocl)
            assert e.last_steps() == o.steps_last()
            assert np.array_equal(e.walks(), o.walks())
            srcs = np.unique(expect_sorted[:, 0] if not remove_dups else expect_sorted[expect_sorted[:, 0] != expect_sorted[:, 1]][:, 0])
            assert np.array_equal(e.touched_vertices(), srcs)


def test_deletes_creating_isolated_vertices(wharf, wo):
    """Deleting every edge of some vertices: walkers standing there stop (documented sink policy) and resume
    when an edge comes back."""
    n = 64
    off, adj = wo.rmat_graph(6, 200, seed=5)
    o = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=9)
    with make(wharf, n, off, adj, walks_per_vertex=3, walk_length=9) as e:
        o.generate(), e.generate_initial_random_walks()
        row_ptr, col = o.csr()
        victims = [int(v) for v in np.argsort(np.diff(row_ptr.astype(np.int64)))[::-1][:3]]
        b = np.array([[v, c] for v in victims for c in col[row_ptr[v]:row_ptr[v + 1]]] , dtype=np.uint32)
        b = np.concatenate([b, b[:, ::-1]])
        o.update(b, delete=True, remove_dups=True)
        e.delete_edges_batch(b.copy(), remove_dups=True)
        assert all(e.degree(v) == 0 for v in victims)
        assert np.array_equal(e.walks(), o.walks())
        assert (o.walks() == wo.NO_VERTEX).any()
        assert e.last_steps() == o.steps_last()
        back = np.array([[victims[0], 1], [1, victims[0]]], dtype=np.uint32)
        o.update(back, delete=False, remove_dups=True)
        e.insert_edges_batch(back.copy(), remove_dups=True)
        assert np.array_equal(e.walks(), o.walks())
        assert np.array_equal(e.affected_walks()[0], o.affected()[0])


def test_empty_graph_and_empty_batches(wharf):
    with wharf.WharfMH(10, config=wharf.Config(walks_per_vertex=2, walk_length=5)) as e:   # WharfMH(n, m): wharfmh.h:26
        assert e.number_of_vertices() == 10 and e.number_of_edges() == 0
        assert e.generate_initial_random_walks() == 0
        w = e.walks()
        assert np.array_equal(w[:, 0], np.tile(np.arange(10, dtype=np.uint32), 2))
        assert (w[:, 1:] == wharf.NO_VERTEX).all()
        assert e.walk(13) == "3 "
        assert e.insert_edges_batch(np.empty((0, 2), dtype=np.uint32)).shape[0] == 0
        affected = e.insert_edges_batch(np.array([[3, 4], [4, 3]], dtype=np.uint32), remove_dups=True)
        assert affected.tolist() == [3, 4, 13, 14]
        assert e.walk(3) == "3 4 3 4 3 "
        with pytest.raises(wharf.WharfError):
            e.insert_edges_batch(np.array([[3, 10]], dtype=np.uint32))       # id >= n
        assert e.insert_edges_batch(np.array([[7, 7]], dtype=np.uint32), remove_dups=True).shape[0] == 0   # only a self-loop


def test_constructor_refuses_a_broken_csr(wharf):
    """graph/wharfmh.h:58-103 takes rows that are sorted and unique with ids < n; anything else would send the walk
    kernels out of bounds or break the sorted-set merges, so the constructor refuses it (WMH_ERR_INVALID)."""
    off = np.array([0, 2, 3, 5], dtype=np.uint64)
    good = np.array([1, 2, 0, 0, 1], dtype=np.uint32)
    with wharf.WharfMH(4, 5, off, good, config=wharf.Config(walks_per_vertex=1, walk_length=3)) as e:
        assert e.number_of_edges() == 5
    for bad_off, bad_adj, what in [
        (off, np.array([1, 2, 0, 0, 9], dtype=np.uint32), "neighbour id"),            # id >= n
        (off, np.array([2, 1, 0, 0, 1], dtype=np.uint32), "ascending"),               # unsorted row
        (off, np.array([1, 1, 0, 0, 1], dtype=np.uint32), "ascending"),               # duplicate
        (np.array([0, 3, 2, 5], dtype=np.uint64), good, "offsets"),                   # not monotone
        (np.array([0, 2, 3, 7], dtype=np.uint64), good, "offsets"),                   # beyond m
    ]:
        with pytest.raises(wharf.WharfError) as err:
            wharf.WharfMH(4, 5, bad_off, bad_adj, config=wharf.Config(walks_per_vertex=1, walk_length=3))
        assert err.value.code == wharf._capi.ERR_INVALID and what in str(err.value)
    # a self-loop and an empty last row are fine
    with wharf.WharfMH(3, 2, np.array([0, 1, 2], dtype=np.uint64), np.array([0, 2], dtype=np.uint32),
                       config=wharf.Config(walks_per_vertex=1, walk_length=3)) as e:
        assert e.generate_initial_random_walks() == 3 + 1          # vertex 0 loops (3 steps), 1 -> 2 (sink), 2 isolated


def test_retrieval_api(wharf, wo):
    off, adj = wo.rmat_graph(8, 1500)
    o = wo.Oracle(256, off, adj, walks_per_vertex=3, walk_length=12)
    o.generate()
    ref = o.walks()
    with make(wharf, 256, off, adj, walks_per_vertex=3, walk_length=12) as e:
        e.generate_initial_random_walks()
        for w in (0, 5, 255, 256, 767):
            row = ref[w][ref[w] != wo.NO_VERTEX]
            assert e.walk(w) == "".join(f"{v} " for v in row)
            assert np.array_equal(e.walk_vertices(w), row)
            for p in range(len(row)):
                assert e.vertex_at_walk(w, p) == row[p]
        assert np.array_equal(e.walks(100, 37), ref[100:137])
        assert np.array_equal(e.walks_gather([700, 3, 3, 90]), ref[[700, 3, 3, 90]])
        with pytest.raises(wharf.WharfError):
            e.walk(768)
        # corpus fingerprint = wrapping sum of per-walk FNV-1a of the text form
        assert e.corpus_fingerprint() == wo.corpus_fingerprint(ref)
        fnv = 0xcbf29ce484222325                                     # the same by hand for walk 5
        for ch in "".join(f"{v} " for v in ref[5][ref[5] != wo.NO_VERTEX]).encode():
            fnv = ((fnv ^ ch) * 0x100000001b3) & ((1 << 64) - 1)
        assert wo.corpus_fingerprint(ref[5:6]) == fnv
        mem = e.memory_footprint()
        assert mem["walks_bytes"] == 768 * 12 * 4 and mem["total_bytes"] > mem["walks_bytes"]
        e.destroy_index()
        with pytest.raises(wharf.WharfError):
            e.walk(0)
        e.set_config(walks_per_vertex=2, walk_length=6)
        o2 = wo.Oracle(256, off, adj, walks_per_vertex=2, walk_length=6)
        o2.generate()
        e.generate_initial_random_walks()
        assert np.array_equal(e.walks(), o2.walks())


def test_inverted_index_find_next(wharf, wo):
    """InvertedIndex::find_next semantics (walks/inverted_index.h:74-92): the entry (walk*L+pos -> next) is stored
    at the vertex the walker stands on; the last position stores the sentinel."""
    off, adj = wo.rmat_graph(7, 600)
    L = 7
    o = wo.Oracle(128, off, adj, walks_per_vertex=2, walk_length=L)
    o.generate()
    ref = o.walks()
    with make(wharf, 128, off, adj, walks_per_vertex=2, walk_length=L) as e:
        e.generate_initial_random_walks()
        e.build_index()
        sizes = np.bincount(ref[ref != wo.NO_VERTEX].astype(np.int64), minlength=128)
        for v in (0, 1, 17, 127):
            assert e.index_size(v) == sizes[v]
        for w in (0, 3, 130, 255):
            for p in range(L):
                v = ref[w, p]
                if v == wo.NO_VERTEX:
                    break
                nxt, found = e.find_next(int(v), w, p)
                expect = wharf.REF_SENTINEL if (p == L - 1 or ref[w, p + 1] == wo.NO_VERTEX) else int(ref[w, p + 1])
                assert found and nxt == expect
                other = (int(v) + 1) % 128
                if ref[w, p] != other:
                    assert not e.find_next(other, w, p)[1]


def test_sharded_corpus_equals_global(wharf, wo):
    """Walks shard by start vertex (wmh_walk_owner) with the graph replicated; the union of the shards
    must be the single-engine corpus, before and after batches."""
    n = 1 << 9
    off, adj = wo.rmat_graph(9, 3000)
    o = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=11)
    o.generate()
    batches = [wo.generate_batch(30, n, s, False, False) for s in range(2)]
    for world in (2, 3):
        engines = [make(wharf, n, off, adj, walks_per_vertex=3, walk_length=11, rank=r, world=world,
                        search_mode=wharf.SEARCH_INDEX if r == 0 else wharf.SEARCH_SCAN) for r in range(world)]
        oo = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=11)
        oo.generate()
        steps = sum(e.generate_initial_random_walks() for e in engines)
        assert steps == oo.steps_last()
        for k in range(3):
            full = np.empty_like(oo.walks())
            for e in engines:
                full[e.local_walk_ids().astype(np.int64)] = e.walks()
            assert np.array_equal(full, oo.walks())
            if k == 2:
                break
            oo.update(batches[k], delete=False, remove_dups=True)
            ids = np.sort(np.concatenate([e.insert_edges_batch(batches[k].copy(), remove_dups=True) for e in engines]))
            assert np.array_equal(ids, oo.affected()[0])
        for e in engines:
            e.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("n,world,wpv", [(512, 2, 2), (509, 3, 2), (1000, 8, 10), (7, 8, 3), (64, 5, 10), (300, 4, 10), (129, 2, 5)])
def test_shard_rule_is_a_partition(wharf, n, world, wpv):
    """Every walk has exactly one owner: whole replicas first (wpv // world per rank); the remaining replicas go to
    gcd(world, rest) groups of ranks, sharded by start vertex inside a group (one vertex per block of group-size ids and
    rank, rotated per block).  Local order = ascending global id, consistent with wmh_walk_owner and the maps."""
    import math
    off = np.zeros(n, dtype=np.uint64)
    adj = np.zeros(0, dtype=np.uint32)
    owner = np.full(n * wpv, -1, dtype=np.int64)
    q = wpv // world
    for r in range(world):
        with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=3, rank=r, world=world) as e:
            ids = e.local_walk_ids().astype(np.int64)
            rest = wpv - world * q
            g = math.gcd(world, rest) if rest else world
            sub_world, shared = world // g, rest // g
            sub_rank = r % sub_world
            starts_per_shared = (n - sub_rank + sub_world - 1) // sub_world if sub_rank < n else 0
            assert ids.shape[0] == e.number_of_local_walks() == q * n + shared * starts_per_shared
            assert (np.diff(ids) > 0).all()
            assert (owner[ids] == -1).all()
            owner[ids] = r
            whole = ids[: q * n]
            assert np.array_equal(whole, np.arange(r * q * n, (r + 1) * q * n))
            assert all(e.walk_owner(int(g)) == r for g in ids[:: max(1, ids.shape[0] // 64)])
            e.generate_initial_random_walks()
            assert np.array_equal(e.local_walk_ids().astype(np.int64), ids)
            assert all(e.global_to_local_walk(int(g)) == i for i, g in list(enumerate(ids))[:: max(1, ids.shape[0] // 64)])
            assert np.array_equal(e.walks()[:, 0], (ids % n).astype(np.uint32))
    assert (owner >= 0).all()
    if world == 8 and n == 1000:
        shared = owner[8 * n: 9 * n]                                   # replica 8 is sharded among ranks 0..3
        assert set(shared.tolist()) == {0, 1, 2, 3} and set(owner[9 * n:].tolist()) == {4, 5, 6, 7}
        assert len({int(shared[b * 4]) for b in range(100)}) > 1       # rotated from block to block


@pytest.mark.gpu
def test_pairings_device(wharf, golden):
    import ctypes as C
    import torch
    lib = wharf._capi.load()
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(3)
    x = rng.integers(0, 2**32, 100000, dtype=np.uint64)
    y = rng.integers(0, 2**32, 100000, dtype=np.uint64)
    x[:4] = [2**32 - 1, 0, 5, 65535]
    y[:4] = [2**32 - 1, 0, 2**31, 65535]
    tx, ty = (torch.from_numpy(a.view(np.int64)).to(dev) for a in (x, y))
    tz, ux, uy = (torch.empty_like(tx) for _ in range(3))
    over = C.c_uint64()
    wharf._capi.check(lib.wmh_szudzik_pair_u64(0, tx.data_ptr(), ty.data_ptr(), tz.data_ptr(), x.shape[0], C.byref(over)))
    assert over.value == 0
    z = tz.cpu().numpy().view(np.uint64)
    big = np.maximum(x, y).astype(object)
    expect = np.array([int(b) * int(b) + int(b) + int(a) if int(yy) >= int(xx) else int(xx) * int(xx) + int(yy)
                       for a, b, xx, yy in zip(x, big, x, y)], dtype=object)
    assert all(int(a) == int(b) for a, b in zip(z[:2000], expect[:2000]))
    wharf._capi.check(lib.wmh_szudzik_unpair_u64(0, tz.data_ptr(), ux.data_ptr(), uy.data_ptr(), x.shape[0]))
    assert np.array_equal(ux.cpu().numpy().view(np.uint64), x) and np.array_equal(uy.cpu().numpy().view(np.uint64), y)
    for a, b, c in golden["pair64"]:
        assert wharf.szudzik_pair(a, b) == c and wharf.szudzik_unpair(c) == (a, b)
    for zz, a, b in golden["unpair64"]:
        assert wharf.szudzik_unpair(zz) == (a, b)
    # overflow is reported, not silently wrapped
    tx[0] = 2**32
    wharf._capi.check(lib.wmh_szudzik_pair_u64(0, tx.data_ptr(), ty.data_ptr(), tz.data_ptr(), 10, C.byref(over)))
    assert over.value == 1
    # u32 codes: tests/pairings.cpp:44-73 domain
    a32 = torch.from_numpy(rng.integers(0, 65536, 50000).astype(np.int32)).to(dev)
    b32 = torch.from_numpy(rng.integers(0, 65536, 50000).astype(np.int32)).to(dev)
    z32, ua, ub = (torch.empty_like(a32) for _ in range(3))
    wharf._capi.check(lib.wmh_szudzik_pair_u32(0, a32.data_ptr(), b32.data_ptr(), z32.data_ptr(), 50000, C.byref(over)))
    wharf._capi.check(lib.wmh_szudzik_unpair_u32(0, z32.data_ptr(), ua.data_ptr(), ub.data_ptr(), 50000))
    assert over.value == 0 and torch.equal(ua, a32) and torch.equal(ub, b32)
    # triplets: tests/pairings.cpp:95-126 domain
    ta = torch.from_numpy(rng.integers(0, 65536, 50000).astype(np.int64)).to(dev)
    tb = torch.from_numpy(rng.integers(0, 65536, 50000).astype(np.int64)).to(dev)
    tc = torch.from_numpy(rng.integers(0, 2**32 - 1, 50000).astype(np.int64)).to(dev)
    tt, ra, rb, rc = (torch.empty_like(ta) for _ in range(4))
    wharf._capi.check(lib.wmh_szudzik_pair_triplet_u64(0, ta.data_ptr(), tb.data_ptr(), tc.data_ptr(), tt.data_ptr(), 50000, C.byref(over)))
    wharf._capi.check(lib.wmh_szudzik_unpair_triplet_u64(0, tt.data_ptr(), ra.data_ptr(), rb.data_ptr(), rc.data_ptr(), 50000))
    assert over.value == 0 and torch.equal(ra, ta) and torch.equal(rb, tb) and torch.equal(rc, tc)
    # known answers made by the unmodified reference (tests/golden/make_golden.py): the DEVICE ops against them, not
    # against their own inverse - pair<u32>, pair_triplet<u64>, unpair_triplet<u64> (tests/pairings.cpp:29-37,75-90)
    def as_i32(v):
        return torch.from_numpy(np.array(v, dtype=np.uint32).view(np.int32)).to(dev)

    def as_i64(v):
        return torch.from_numpy(np.array(v, dtype=np.uint64).view(np.int64)).to(dev)

    g32 = golden["pair32"]
    k = len(g32)
    gx, gy, gz = as_i32([r[0] for r in g32]), as_i32([r[1] for r in g32]), as_i32([r[2] for r in g32])
    oz, ox, oy = (torch.empty_like(gx) for _ in range(3))
    wharf._capi.check(lib.wmh_szudzik_pair_u32(0, gx.data_ptr(), gy.data_ptr(), oz.data_ptr(), k, C.byref(over)))
    assert over.value == 0 and torch.equal(oz, gz)
    wharf._capi.check(lib.wmh_szudzik_unpair_u32(0, gz.data_ptr(), ox.data_ptr(), oy.data_ptr(), k))
    assert torch.equal(ox, gx) and torch.equal(oy, gy)
    g64 = golden["triplet64"] + [r for r in golden["triplet32"] if r not in golden["triplet64"]]
    k = len(g64)
    ga, gb, gc, gt = (as_i64([r[i] for r in g64]) for i in range(4))
    ot, oa, ob, oc = (torch.empty_like(ga) for _ in range(4))
    wharf._capi.check(lib.wmh_szudzik_pair_triplet_u64(0, ga.data_ptr(), gb.data_ptr(), gc.data_ptr(), ot.data_ptr(), k, C.byref(over)))
    assert over.value == 0 and torch.equal(ot, gt)
    assert [123, 25, 200, 229643916] in g64                       # the reference's own triplet KAT
    wharf._capi.check(lib.wmh_szudzik_unpair_triplet_u64(0, gt.data_ptr(), oa.data_ptr(), ob.data_ptr(), oc.data_ptr(), k))
    assert torch.equal(oa, ga) and torch.equal(ob, gb) and torch.equal(oc, gc)
    g64p = golden["pair64"]
    k = len(g64p)
    px, py, pz = (as_i64([r[i] for r in g64p]) for i in range(3))
    qz, qx, qy = (torch.empty_like(px) for _ in range(3))
    wharf._capi.check(lib.wmh_szudzik_pair_u64(0, px.data_ptr(), py.data_ptr(), qz.data_ptr(), k, C.byref(over)))
    assert over.value == 0 and torch.equal(qz, pz)
    uz = as_i64([r[0] for r in golden["unpair64"]])
    rx, ry = torch.empty_like(uz), torch.empty_like(uz)
    wharf._capi.check(lib.wmh_szudzik_unpair_u64(0, uz.data_ptr(), rx.data_ptr(), ry.data_ptr(), uz.shape[0]))
    assert rx.cpu().numpy().view(np.uint64).tolist() == [r[1] for r in golden["unpair64"]]
    assert ry.cpu().numpy().view(np.uint64).tolist() == [r[2] for r in golden["unpair64"]]


def test_compressed_walks(wharf, wo):
    """walks/compressed_walks.h: per-vertex ascending codes pair(walk_id*L+pos, next) and find_next over them."""
    import ctypes as C
    off, adj = wo.rmat_graph(7, 600)
    L = 6
    o = wo.Oracle(128, off, adj, walks_per_vertex=2, walk_length=L)
    o.generate()
    ref = o.walks()
    with make(wharf, 128, off, adj, walks_per_vertex=2, walk_length=L) as e:
        e.generate_initial_random_walks()
        ptr, codes = e.compressed_walks()
        expect = [[] for _ in range(128)]
        for w in range(256):
            for p in range(L):
                v = ref[w, p]
                if v == wo.NO_VERTEX:
                    break
                nxt = 0xFFFFFFFE if (p == L - 1 or ref[w, p + 1] == wo.NO_VERTEX) else int(ref[w, p + 1])
                expect[v].append(wo.lib().wo_pair64(w * L + p, nxt))
        for v in range(128):
            assert codes[ptr[v]:ptr[v + 1]].tolist() == sorted(expect[v])
        for w, p in [(0, 0), (5, 3), (200, 5)]:
            v = int(ref[w, p])
            seg = np.ascontiguousarray(codes[ptr[v]:ptr[v + 1]])
            want = wo.lib().wo_compressed_find_next(C.c_void_p(seg.ctypes.data), seg.shape[0], w, p, L)
            assert e.compressed_find_next(v, w, p) == (want, True)
        assert e.compressed_find_next(int(ref[0, 0]), 1, 1)[1] in (False, True)


def test_rmat_generator_device(wharf, wo, golden):
    import ctypes as C
    import torch
    lib = wharf._capi.load()
    for g in golden["genbatch"]:
        total = g["edges"] * (1 if g["directed"] else 2)
        out = torch.empty(2 * total, dtype=torch.int32, device="cuda:0")
        k = C.c_uint64()
        wharf._capi.check(lib.wmh_generate_batch_of_edges_device(0, g["edges"], g["vertices"], g["seed"], g["self_loops"], g["directed"],
                                                                 0.5, 0.2, 0.1, out.data_ptr(), C.byref(k)))
        pairs = out[: 2 * k.value].cpu().numpy().view(np.uint32).reshape(-1, 2)
        assert k.value == g["count"] and sha(pairs) == g["sha256"]
    # CSR from pairs
    n = 1 << 10
    pairs = wo.generate_batch(5000, n + 1, 42, False, False)
    off, adj = wo.csr_from_sorted_pairs(n, pairs)
    tp = torch.from_numpy(pairs.view(np.int32).reshape(-1)).cuda()
    toff = torch.empty(n + 1, dtype=torch.int64, device="cuda:0")
    tadj = torch.empty(pairs.shape[0], dtype=torch.int32, device="cuda:0")
    wharf._capi.check(lib.wmh_pairs_to_csr_device(0, n, tp.data_ptr(), pairs.shape[0], toff.data_ptr(), tadj.data_ptr()))
    assert np.array_equal(toff.cpu().numpy()[:n].astype(np.uint64), off) and toff[n].item() == pairs.shape[0]
    assert np.array_equal(tadj.cpu().numpy().view(np.uint32), adj)
    # engine straight from device arrays
    cfg = wharf.Config(walks_per_vertex=2, walk_length=10)
    e = wharf.WharfMH(n, pairs.shape[0], toff.data_ptr(), tadj.data_ptr(), config=cfg, device_arrays=True)
    o = wo.Oracle(n, off, adj, walks_per_vertex=2, walk_length=10)
    assert e.generate_initial_random_walks() == o.generate()
    assert np.array_equal(e.walks(), o.walks())
    e.destroy()


@pytest.mark.parametrize("model,init", [("deepwalk", "weight"), ("node2vec", "weight"), ("node2vec", "random"), ("node2vec", "burnin")])
def test_keyed_mh_mode_vs_oracle(wharf, wo, model, init):
    """Metropolis-Hastings mode (deterministic_mode = false) with keyed draws: bit-exact against the keyed CPU
    restatement, through initial generation and symmetric insert/delete batches."""
    n = 1 << 9
    off, adj = wo.rmat_graph(9, 4000, seed=11)
    m = wo.NODE2VEC if model == "node2vec" else wo.DEEPWALK
    ini = {"weight": wo.INIT_WEIGHT, "random": wo.INIT_RANDOM, "burnin": wo.INIT_BURNIN}[init]
    o = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=15, model=m, p=0.5, q=2.0, deterministic=False, init_strategy=ini, seed=77)
    with make(wharf, n, off, adj, walks_per_vertex=3, walk_length=15, random_walk_model=m, paramP=0.5, paramQ=2.0,
              deterministic_mode=False, sampler_init_strategy=ini, seed=77) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert np.array_equal(e.walks(), o.walks())
        for s, delete in [(0, False), (1, False), (0, True), (2, False)]:
            b = wo.generate_batch(25, n, s, False, False)
            o.update(b, delete=delete, remove_dups=True)
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert np.array_equal(e.affected_walks()[0], o.affected()[0])
            assert np.array_equal(e.walks(), o.walks())
    # a different seed gives a different corpus
    o2 = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=15, model=m, p=0.5, q=2.0, deterministic=False, init_strategy=ini, seed=78)
    o2.generate()
    assert not np.array_equal(o2.walks(), o.walks())


@pytest.mark.parametrize("init", ["weight", "random", "burnin"])
@pytest.mark.parametrize("eager", ["1", "0"])
def test_bulk_sampler_init_equals_first_touch(wharf, wo, init, eager, monkeypatch):
    """A corpus large enough to touch most sampler states initialises them with the bulk kernel before walking
    (n * wpv * L >= 8 m); WHARF_EAGER_INIT=0 keeps the first-touch path.  Same corpus, bit for bit, as the oracle."""
    monkeypatch.setenv("WHARF_EAGER_INIT", eager)
    n = 1 << 8
    off, adj = wo.rmat_graph(8, 1500, seed=4)
    ini = {"weight": wo.INIT_WEIGHT, "random": wo.INIT_RANDOM, "burnin": wo.INIT_BURNIN}[init]
    kw = dict(walks_per_vertex=10, walk_length=40)
    assert n * 10 * 40 >= 8 * adj.shape[0]
    o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, init_strategy=ini, seed=21, **kw)
    with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False,
              sampler_init_strategy=ini, seed=21, **kw) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert np.array_equal(e.walks(), o.walks())
        b = wo.generate_batch(20, n, 2, False, False)
        o.update(b, delete=False, remove_dups=True)
        e.insert_edges_batch(b.copy(), remove_dups=True)
        assert np.array_equal(e.walks(), o.walks())
        e.destroy_index()
        assert e.generate_initial_random_walks() == o.generate()          # a fresh corpus on the updated graph
        assert np.array_equal(e.walks(), o.walks())


@pytest.mark.parametrize("init", ["weight", "random"])
def test_keyed_mh_directed_batches_vs_oracle(wharf, wo, init):
    """node2vec MH with DIRECTED insert/delete batches (no reverse edges): the sampler cache must stay equal to a
    recomputation from the current graph, walkers may meet sinks, and prev may not be a neighbour of cur."""
    rng = np.random.default_rng(5)
    n = 1 << 8
    off, adj = wo.rmat_graph(8, 1500, seed=4)
    ini = {"weight": wo.INIT_WEIGHT, "random": wo.INIT_RANDOM}[init]
    kw = dict(walks_per_vertex=5, walk_length=20)
    o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.25, q=4.0, deterministic=False, init_strategy=ini, seed=3, **kw)
    with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.25, paramQ=4.0, deterministic_mode=False,
              sampler_init_strategy=ini, seed=3, **kw) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert np.array_equal(e.walks(), o.walks())
        for step in range(8):
            delete = step % 2 == 1
            if delete:
                row_ptr, col = o.csr()
                pick = rng.integers(0, col.shape[0], 60)
                src = np.searchsorted(row_ptr, pick, side="right") - 1
                b = np.stack([src, col[pick]], 1).astype(np.uint32)            # one direction only
            else:
                b = random_batch(rng, n, 40, symmetric=False)
            b = np.ascontiguousarray(b)
            o.update(b, delete=delete, remove_dups=True)
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert np.array_equal(e.affected_walks()[0], o.affected()[0])
            assert e.last_steps() == o.steps_last()
            assert np.array_equal(e.walks(), o.walks()), f"corpus differs after directed batch {step}"


@pytest.mark.parametrize("case", ["rmat10", "rmat12_sparse", "directed_sinks", "short", "many_replicas", "single"])
def test_coalesced_generation_equals_oracle(wharf, wo, case):
    """GENERATE_COALESCED (step-synchronous, walkers of a replica that met advance once) must give the corpus and
    step count of the oracle and of the one-walker-per-thread kernel, including isolated starts, sinks of a
    directed graph, walk_length 1/2 and sharded engines."""
    wpv, L = 10, 80
    if case == "rmat10":
        n = 1 << 10
        off, adj = wo.rmat_graph(10, 8000)
    elif case == "rmat12_sparse":
        n = 1 << 12
        off, adj = wo.rmat_graph(12, 5000, seed=9)            # many isolated vertices
    elif case == "directed_sinks":
        n = 300
        rng = np.random.default_rng(1)
        pairs = np.unique(rng.integers(0, n, (900, 2)).astype(np.uint32), axis=0)
        pairs = pairs[(pairs[:, 0] != pairs[:, 1]) & (pairs[:, 0] % 7 != 0)]     # every 7th vertex is a sink
        off, adj = wo.csr_from_sorted_pairs(n, pairs)
    elif case == "short":
        n, wpv, L = 1 << 9, 3, 2
        off, adj = wo.rmat_graph(9, 3000)
    elif case == "many_replicas":                              # more whole replicas than one fused first-step launch takes (16)
        n, wpv, L = 1 << 9, 37, 6
        off, adj = wo.rmat_graph(9, 2500, seed=4)
    else:
        n, wpv, L = 1 << 9, 1, 1
        off, adj = wo.rmat_graph(9, 3000)
    o = wo.Oracle(n, off, adj, walks_per_vertex=wpv, walk_length=L)
    steps = o.generate()
    with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=L, generate_mode=wharf.GENERATE_COALESCED) as e:
        assert e.generate_initial_random_walks() == steps
        assert np.array_equal(e.walks(), o.walks())
        # and a batch afterwards re-walks from the coalesced corpus exactly like the oracle
        b = wo.generate_batch(20, n, 1, False, False)
        o.update(b, delete=False, remove_dups=True)
        e.insert_edges_batch(b.copy(), remove_dups=True)
        assert np.array_equal(e.walks(), o.walks())
    with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=L, generate_mode=wharf.GENERATE_PER_WALKER) as e:
        assert e.generate_initial_random_walks() == steps
    o2 = wo.Oracle(n, off, adj, walks_per_vertex=wpv, walk_length=L)
    o2.generate()
    for world in (2, 3):
        full = np.empty_like(o2.walks())
        total = 0
        for r in range(world):
            with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=L, rank=r, world=world,
                      generate_mode=wharf.GENERATE_COALESCED) as e:
                total += e.generate_initial_random_walks()
                full[e.local_walk_ids().astype(np.int64)] = e.walks()
        assert total == steps
        assert np.array_equal(full, o2.walks())


@pytest.mark.parametrize("knobs", [{"WHARF_RANK_SCAN": "1"}, {"WHARF_GRAPH": "0"}, {"WHARF_BRANCHES": "1"}, {"WHARF_BRANCHES": "7"}, {"WHARF_FIRST_STEPS": "0"}])
def test_coalesced_schedule_variants(wharf, wo, knobs, monkeypatch):
    """The corpus does not depend on how the coalesced pass is scheduled: library-scan ranking (the path for more than
    134 M vertices), plain launches instead of the captured CUDA graph, one or many parallel graph branches, the first step of
    the whole replicas taken per replica instead of in one sweep."""
    for k, v in knobs.items():
        monkeypatch.setenv(k, v)
    n = 1 << 11
    off, adj = wo.rmat_graph(11, 12000, seed=3)
    o = wo.Oracle(n, off, adj, walks_per_vertex=10, walk_length=80)
    steps = o.generate()
    for rank, world in ((0, 1), (1, 4)):
        with make(wharf, n, off, adj, walks_per_vertex=10, walk_length=80, generate_mode=wharf.GENERATE_COALESCED, rank=rank, world=world) as e:
            for _ in range(2):                                   # the second pass replays the captured graph
                got = e.generate_initial_random_walks()
                assert np.array_equal(e.walks(), o.walks()[e.local_walk_ids().astype(np.int64)])
            if world == 1:
                assert got == steps


@pytest.mark.parametrize("mode", ["det_walker", "det_coalesced", "n2v_mh", "deepwalk_mh", "sharded"])
def test_generate_to_host_equals_generate_then_copy(wharf, wo, mode):
    """wmh_generate_initial_random_walks_to_host (generation overlapped with the export) must deliver the corpus that
    generate + walks() delivers, leave the same corpus on the device, and report the same step count."""
    import torch
    n = 1 << 11
    off, adj = wo.rmat_graph(11, 12000, seed=2)
    kw = dict(walks_per_vertex=7, walk_length=33)
    if mode == "det_walker":
        kw.update(generate_mode=wharf.GENERATE_PER_WALKER)
    elif mode == "det_coalesced":
        kw.update(generate_mode=wharf.GENERATE_COALESCED)
    elif mode == "n2v_mh":
        kw.update(deterministic_mode=False, random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, seed=5)
    elif mode == "deepwalk_mh":
        kw.update(deterministic_mode=False, random_walk_model=wharf.DEEPWALK, seed=5)
    else:
        kw.update(rank=1, world=3, generate_mode=wharf.GENERATE_COALESCED)
    with make(wharf, n, off, adj, **kw) as a, make(wharf, n, off, adj, **kw) as b:
        steps = a.generate_initial_random_walks()
        expect = a.walks()
        out = torch.empty((b.number_of_local_walks(), 33), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
        out[:] = 7
        assert b.generate_initial_random_walks_to_host(out) == steps
        assert np.array_equal(out, expect)
        assert np.array_equal(b.walks(), expect)
        pageable = np.zeros_like(expect)
        assert b.generate_initial_random_walks_to_host(pageable) == steps
        assert np.array_equal(pageable, expect)


def _sym_pairs(pairs):
    pairs = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
    both = np.concatenate([pairs, pairs[:, ::-1]])
    both = both[both[:, 0] != both[:, 1]]
    return np.unique(both, axis=0).astype(np.uint32)


@pytest.mark.parametrize("init", ["weight", "burnin"])
def test_triangle_filter_symmetric_batches_vs_oracle(wharf, wo, init, monkeypatch):
    """node2vec's membership test answered from the aux word of the proposal's own slot (triangle blocks, wmh_neighbour_filter
    == 2) on a triangle-rich symmetric graph: inserts that close triangles between edges of the same batch, a batch that
    deletes triangle edges (bits stay behind: a superset), enough deletes to force a rebuild, then a DIRECTED batch, which
    must drop the engine back to per-row blocks.  The corpus equals the keyed oracle after every step, and equals the
    corpus of an engine that never used triangle blocks (WHARF_TRI=0)."""
    rng = np.random.default_rng(17)
    n = 1 << 8
    off, adj = wo.rmat_graph(8, 3000, seed=9)                         # ~20 neighbours per vertex on 256 vertices: many triangles
    ini = {"weight": wo.INIT_WEIGHT, "burnin": wo.INIT_BURNIN}[init]
    kw = dict(walks_per_vertex=6, walk_length=30)
    o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, init_strategy=ini, seed=5, **kw)
    corpora = []
    with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False,
              sampler_init_strategy=ini, seed=5, **kw) as e:
        assert e.neighbour_filter() == 0
        assert e.generate_initial_random_walks() == o.generate()
        assert e.neighbour_filter() == 2
        assert np.array_equal(e.walks(), o.walks())
        corpora.append(e.walks())
        ops = []
        for step in range(10):
            row_ptr, col = o.csr()
            if step % 3 == 2:                                         # delete existing edges, both directions: triangle edges among them
                src = np.repeat(np.arange(n), np.diff(row_ptr).astype(np.int64))
                pick = rng.choice(col.shape[0], 300 if step != 8 else col.shape[0] // 3, replace=False)
                b, delete = _sym_pairs(np.stack([src[pick], col[pick]], 1)), True
            else:                                                     # a clique on 6 random vertices + random edges: triangles inside one batch
                clique = rng.choice(n, 6, replace=False)
                extra = rng.integers(0, n, (60, 2))
                b, delete = _sym_pairs(np.concatenate([[(u, v) for u in clique for v in clique if u < v], extra])), False
            ops.append((b, delete))
            o.update(b.copy(), delete=delete, remove_dups=True)
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert e.neighbour_filter() == 2
            assert e.filter_check() == 0, f"step {step}: a triangle block lost a member"
            assert np.array_equal(e.affected_walks()[0], o.affected()[0])
            assert np.array_equal(e.walks(), o.walks()), f"step {step}"
            corpora.append(e.walks())
        e.destroy_index()                                             # a fresh corpus on the updated graph, blocks as maintained
        assert e.generate_initial_random_walks() == o.generate() and np.array_equal(e.walks(), o.walks())
        directed = rng.integers(0, n, (40, 2)).astype(np.uint32)      # no reverse edges: the blocks no longer hold
        o.update(directed.copy(), delete=False, remove_dups=True)
        e.insert_edges_batch(directed.copy(), remove_dups=True)
        assert e.neighbour_filter() == 1 and e.filter_check() == 0
        assert np.array_equal(e.walks(), o.walks())
        b = _sym_pairs(rng.integers(0, n, (50, 2)))
        o.update(b.copy(), delete=False, remove_dups=True)
        e.insert_edges_batch(b.copy(), remove_dups=True)
        assert e.neighbour_filter() == 1 and e.filter_check() == 0 and np.array_equal(e.walks(), o.walks())
    monkeypatch.setenv("WHARF_TRI", "0")
    with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False,
              sampler_init_strategy=ini, seed=5, **kw) as e:
        e.generate_initial_random_walks()
        assert e.neighbour_filter() == 1 and np.array_equal(e.walks(), corpora[0])
        for (b, delete), want in zip(ops, corpora[1:]):
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert np.array_equal(e.walks(), want)


@pytest.mark.parametrize("coop", ["0", "1"])
def test_triangle_filter_hub_rows(wharf, wo, coop, monkeypatch):
    """Slots whose BOTH rows are long (hub - hub edges) take the block-per-slot kernels, at build time and when an inserted
    edge joins two hubs; a directed start graph never gets triangle blocks.  coop = 1: membership searches in rows of 1024
    neighbours and more are done by the whole warp (32 probes per round, WHARF_COOP_SEARCH) - the same corpus."""
    monkeypatch.setenv("WHARF_COOP_SEARCH", coop)
    rng = np.random.default_rng(23)
    n, leaves = 6000, 5000
    edges = [(h, v) for h in (0, 1, 2) for v in range(3, leaves)] + [(0, 1)] + rng.integers(3, n, (4000, 2)).tolist()
    pairs = _sym_pairs(edges)
    off, adj = wo.csr_from_sorted_pairs(n, pairs)
    kw = dict(walks_per_vertex=2, walk_length=12)
    o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, seed=8, **kw)
    with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=8, **kw) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert e.neighbour_filter() == 2 and e.degree(0) > 2048 and e.degree(2) > 2048
        assert np.array_equal(e.walks(), o.walks())
        for b, delete in ((_sym_pairs([(0, 2), (1, 2), (5, 6)]), False), (_sym_pairs([(0, 1), (7, 8)]), True), (_sym_pairs([(0, 1)]), False)):
            o.update(b.copy(), delete=delete, remove_dups=True)
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert e.neighbour_filter() == 2 and e.filter_check() == 0 and np.array_equal(e.walks(), o.walks())
    directed = pairs[pairs[:, 0] < pairs[:, 1]]                       # one direction only
    doff, dadj = wo.csr_from_sorted_pairs(n, directed)
    o = wo.Oracle(n, doff, dadj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, seed=8, **kw)
    with make(wharf, n, doff, dadj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=8, **kw) as e:
        assert e.generate_initial_random_walks() == o.generate()
        assert e.neighbour_filter() == 1 and np.array_equal(e.walks(), o.walks())


def test_triangle_filter_symmetry_detection(wharf, wo):
    """The triangle build decides by itself whether the graph is symmetric: a pair is handled once, from its longer row, and
    the count of slots written must reach the edge count.  One missing twin - seen from the longer row or from the shorter
    one - sends the engine to per-row blocks; self-loops are their own twins and keep triangle blocks.  Exact either way."""
    rng = np.random.default_rng(31)
    n = 300
    base = _sym_pairs(np.concatenate([[(0, v) for v in range(1, 120)], rng.integers(0, n, (900, 2))]))
    loops = np.array([(5, 5), (0, 0), (77, 77)], dtype=np.uint32)
    kw = dict(walks_per_vertex=3, walk_length=20)
    cases = {"loops": (np.concatenate([base, loops]), 2)}
    assert ((base[:, 0] == 0) & (base[:, 1] == 50)).sum() == 1
    cases["hub side missing"] = (base[~((base[:, 0] == 0) & (base[:, 1] == 50))], 1)      # 50 -> 0 without 0 -> 50
    cases["leaf side missing"] = (base[~((base[:, 0] == 50) & (base[:, 1] == 0))], 1)     # 0 -> 50 without 50 -> 0
    for name, (pairs, want) in cases.items():
        pairs = np.unique(pairs.astype(np.int64), axis=0).astype(np.uint32)
        off, adj = wo.csr_from_sorted_pairs(n, pairs)
        o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, seed=3, **kw)
        with make(wharf, n, off, adj, random_walk_model=wo.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=3, **kw) as e:
            assert e.generate_initial_random_walks() == o.generate(), name
            assert e.neighbour_filter() == want and e.filter_check() == 0, name
            assert np.array_equal(e.walks(), o.walks()), name
            b = _sym_pairs(rng.integers(0, n, (30, 2)))
            o.update(b.copy(), delete=False, remove_dups=True)
            e.insert_edges_batch(b.copy(), remove_dups=True)
            assert e.neighbour_filter() == want and e.filter_check() == 0 and np.array_equal(e.walks(), o.walks()), name
        o.close()


@pytest.mark.parametrize("positions_per_chunk", [1, 4, 7])
def test_index_in_chunks_of_positions(wharf, wo, positions_per_chunk, monkeypatch):
    """The inverted index keeps 32-bit cell ids at any corpus size by sorting the corpus in chunks of whole positions
    (WHARF_INDEX_CHUNK_CELLS forces the chunking a corpus of more than 2^31 cells gets): entry counts, find_next at every
    position incl. the chunk borders, the INDEX search (affected ids, earliest positions) and the corpora after batches equal
    the oracle's; the paired corpus (CompressedWalks) asks for a one-chunk index and says so."""
    rng = np.random.default_rng(5)
    n, wpv, L = 1 << 9, 3, 19
    off, adj = wo.rmat_graph(9, 5 * n, seed=2)
    monkeypatch.setenv("WHARF_INDEX_CHUNK_CELLS", str(positions_per_chunk * n * wpv))
    o = wo.Oracle(n, off, adj, walks_per_vertex=wpv, walk_length=L)
    with make(wharf, n, off, adj, walks_per_vertex=wpv, walk_length=L, search_mode=wharf.SEARCH_INDEX) as e:
        assert e.generate_initial_random_walks() == o.generate()
        ref = o.walks()
        sizes = np.bincount(ref[ref != wo.NO_VERTEX].astype(np.int64), minlength=n)
        assert np.array_equal(e.index_sizes(), sizes)
        for v in (0, 1, 200, n - 1):
            assert e.index_size(v) == sizes[v]
        for w in (0, 5, n + 3, n * wpv - 1):
            for p in range(L):
                v = ref[w, p]
                if v == wo.NO_VERTEX:
                    break
                nxt, found = e.find_next(int(v), w, p)
                expect = wharf.REF_SENTINEL if (p == L - 1 or ref[w, p + 1] == wo.NO_VERTEX) else int(ref[w, p + 1])
                assert found and nxt == expect, (w, p)
                assert not e.find_next((int(v) + 1) % n, w, p)[1] or ref[w, p] == (int(v) + 1) % n
        for step in range(4):
            b = random_batch(rng, n, 30 + 10 * step)
            delete = step == 2
            o.update(b.copy(), delete=delete, remove_dups=True)
            (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
            assert e.last_search_mode() == wharf.SEARCH_INDEX
            ids, pos = e.affected_walks()
            assert np.array_equal(ids, o.affected()[0]) and np.array_equal(pos, o.affected()[1])
            assert np.array_equal(e.walks(), o.walks())
        ref = o.walks()
        assert np.array_equal(e.index_sizes(), np.bincount(ref[ref != wo.NO_VERTEX].astype(np.int64), minlength=n))
        with pytest.raises(wharf.WharfError):
            e.compressed_walks()
