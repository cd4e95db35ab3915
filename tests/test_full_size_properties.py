"""BASELINE.json's full-size configuration (configs[1]: RMAT scale-20, 1 M vertices, 16.7 M edges, 10 x 80 walks) through
size-independent properties - the CPU oracle cannot produce 839 M steps in test time, so at this size the checks are:

  * a checksum of checksums: the corpus fingerprint is a wrapping sum of per-walk hashes, so the fingerprints of the
    shards of a sharded run must add up to the fingerprint of the unsharded corpus (deterministic AND keyed MH mode:
    draws are keyed by the global walk id);
  * schedule independence: the coalesced schedule and the one-walker-per-thread kernel produce the same corpus;
  * walk validity: every stored step (v[p], v[p+1]) is an edge of the graph, every walk starts at walk_id % n, a walk
    stops only on a vertex without neighbours;
  * reproducibility of keyed MH (same seed -> same corpus, another seed -> another corpus);
  * insert-then-delete round trip: the adjacency ends as the original edge set minus the batch, bit for bit, both
    updates report the same affected walks, and the corpus stays a set of walks of the updated graph.
Inputs come from the reference's own generator restated on the device (byte-identical: test_rmat_generator_device).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCALE, EDGES_HALF, WPV, LEN = 20, 8 << 20, 10, 80


@pytest.fixture(scope="module")
def graph(wharf):
    import torch
    lib = wharf._capi.load()
    n, vertices = 1 << SCALE, (1 << SCALE) + 1
    dev = torch.device("cuda:0")
    pairs = torch.empty(4 * EDGES_HALF, dtype=torch.int32, device=dev)
    k = C.c_uint64()
    wharf._capi.check(lib.wmh_generate_batch_of_edges_device(0, EDGES_HALF, vertices, 42, 0, 0, 0.5, 0.2, 0.1, pairs.data_ptr(), C.byref(k)))
    m = int(k.value)
    offsets = torch.empty(n + 1, dtype=torch.int64, device=dev)
    edges = torch.empty(m, dtype=torch.int32, device=dev)
    wharf._capi.check(lib.wmh_pairs_to_csr_device(0, n, pairs.data_ptr(), m, offsets.data_ptr(), edges.data_ptr()))
    del pairs
    keys = (torch.repeat_interleave(torch.arange(n, device=dev, dtype=torch.int64), offsets[1:] - offsets[:-1]) << 32) | (edges.to(torch.int64) & 0xFFFFFFFF)
    return {"n": n, "m": m, "offsets": offsets, "edges": edges, "keys": keys, "degree": offsets[1:] - offsets[:-1]}


def engine(wharf, g, **kw):
    cfg = wharf.Config(walks_per_vertex=WPV, walk_length=LEN, search_mode=wharf.SEARCH_SCAN, **kw)
    return wharf.WharfMH(g["n"], g["m"], g["offsets"].data_ptr(), g["edges"].data_ptr(), config=cfg, device_arrays=True)


def device_corpus(e):
    """Position-major walk matrix [L][W_local] as a torch view of the engine's own buffer (no copy)."""
    import torch

    class View:                                          # __cuda_array_interface__ around the raw device pointer
        __cuda_array_interface__ = {"data": (e.device_walks_ptr(), False), "shape": (LEN, e.number_of_local_walks()), "typestr": "<i4",
                                    "version": 3, "strides": None}
    return torch.as_tensor(View(), device="cuda:0")


def check_walks_are_valid(g, e, keys=None, degree=None):
    import torch
    keys = g["keys"] if keys is None else keys
    degree = g["degree"] if degree is None else degree
    walks = device_corpus(e).to(torch.int64) & 0xFFFFFFFF
    n, w_local = g["n"], walks.shape[1]
    starts = torch.from_numpy((e.local_walk_ids() % np.uint64(g["n"])).astype(np.int64)).to(walks.device)
    assert torch.equal(walks[0], starts), "a walk does not start at walk_id % n"
    bad = 0
    for p in range(LEN - 1):
        a, b = walks[p], walks[p + 1]
        moved = b != 0xFFFFFFFF
        alive = a != 0xFFFFFFFF
        assert not bool((moved & ~alive).any()), "a stopped walker moved again"
        q = (a[moved] << 32) | b[moved]
        pos = torch.searchsorted(keys, q).clamp(max=keys.shape[0] - 1)
        bad += int((keys[pos] != q).sum().item())
        stopped_here = alive & ~moved
        assert not bool((degree[a[stopped_here]] != 0).any()), "a walker stopped on a vertex that has neighbours"
    assert bad == 0, f"{bad} stored steps are not edges of the graph"
    assert n * WPV >= w_local


def test_deterministic_schedules_and_shards_agree(wharf, graph):
    fps = {}
    for name, mode in (("per_walker", wharf.GENERATE_PER_WALKER), ("coalesced", wharf.GENERATE_COALESCED)):
        with engine(wharf, graph, deterministic_mode=True, generate_mode=mode) as e:
            steps = e.generate_initial_random_walks()
            fps[name] = (e.corpus_fingerprint(), steps)
            if name == "coalesced":
                check_walks_are_valid(graph, e)
    assert fps["per_walker"] == fps["coalesced"]
    for world in (3, 8):
        total_fp, total_steps = 0, 0
        for r in range(world):
            with engine(wharf, graph, deterministic_mode=True, generate_mode=wharf.GENERATE_COALESCED, rank=r, world=world) as e:
                total_steps += e.generate_initial_random_walks()
                total_fp = (total_fp + e.corpus_fingerprint()) & 0xFFFFFFFFFFFFFFFF
        assert (total_fp, total_steps) == fps["coalesced"], f"shards of world {world} do not add up to the unsharded corpus"


def test_keyed_mh_full_size(wharf, graph):
    kw = dict(random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False)
    with engine(wharf, graph, seed=1, **kw) as e:
        steps = e.generate_initial_random_walks()
        fp = e.corpus_fingerprint()
        check_walks_are_valid(graph, e)
        e.destroy_index()
        assert e.generate_initial_random_walks() == steps and e.corpus_fingerprint() == fp, "keyed MH is not reproducible"
    assert steps == int((graph["degree"] > 0).sum().item()) * WPV * LEN      # a symmetric graph has no sinks
    with engine(wharf, graph, seed=2, **kw) as e:
        e.generate_initial_random_walks()
        assert e.corpus_fingerprint() != fp
    total_fp, total_steps = 0, 0
    for r in range(4):
        with engine(wharf, graph, seed=1, rank=r, world=4, **kw) as e:
            total_steps += e.generate_initial_random_walks()
            total_fp = (total_fp + e.corpus_fingerprint()) & 0xFFFFFFFFFFFFFFFF
    assert (total_fp, total_steps) == (fp, steps), "MH shards do not add up to the unsharded corpus"


def test_insert_then_delete_round_trip(wharf, graph):
    import torch
    lib = wharf._capi.load()
    batch = torch.empty(4 * 100000, dtype=torch.int32, device="cuda:0")
    k = C.c_uint64()
    wharf._capi.check(lib.wmh_generate_batch_of_edges_device(0, 100000, graph["n"], 7, 0, 0, 0.5, 0.2, 0.1, batch.data_ptr(), C.byref(k)))
    pairs = int(k.value)
    row_ptr0 = graph["offsets"].cpu().numpy().astype(np.uint64)
    col0 = graph["edges"].cpu().numpy().view(np.uint32)
    with engine(wharf, graph, random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False, seed=3) as e:
        e.generate_initial_random_walks()
        torch.cuda.synchronize()
        a_ins = e.insert_edges_batch_device(batch.data_ptr(), pairs, remove_dups=True)
        ids_ins = e.affected_walks()[0]
        assert e.number_of_edges() > graph["m"]
        # the corpus is a set of walks of the UPDATED graph
        rp, col = e.flatten_graph()
        keys = (torch.repeat_interleave(torch.arange(graph["n"], dtype=torch.int64), torch.from_numpy(np.diff(rp.astype(np.int64)))) << 32) \
            | torch.from_numpy(col.astype(np.int64))
        check_walks_are_valid(graph, e, keys=keys.cuda(), degree=torch.from_numpy(np.diff(rp.astype(np.int64))).cuda())
        a_del = e.delete_edges_batch_device(batch.data_ptr(), pairs, remove_dups=True)
        ids_del = e.affected_walks()[0]
        assert a_ins == a_del and np.array_equal(ids_ins, ids_del)
        rp, col = e.flatten_graph()
        # edges of the batch that already existed are gone too: compare with the original minus the batch
        b = batch[: 2 * pairs].cpu().numpy().view(np.uint32).reshape(-1, 2).astype(np.int64)
        orig = (np.repeat(np.arange(graph["n"], dtype=np.int64), np.diff(row_ptr0.astype(np.int64))) << 32) | col0.astype(np.int64)
        want = np.setdiff1d(orig, (b[:, 0] << 32) | b[:, 1], assume_unique=False)
        have = (np.repeat(np.arange(graph["n"], dtype=np.int64), np.diff(rp.astype(np.int64))) << 32) | col.astype(np.int64)
        assert np.array_equal(have, want), "adjacency after insert + delete is not original minus batch"
        check_walks_are_valid(graph, e, keys=torch.from_numpy(have).cuda(), degree=torch.from_numpy(np.diff(rp.astype(np.int64))).cuda())


def test_walk_keys_beyond_32_bits_equal_the_oracle(wharf):
    """walk_id * walk_length + position is a uint32_t in the reference (walks/inverted_index.h:14,76; graph/wharfmh.h:316) and
    wraps beyond 2^32 - the engine's keys and cell offsets are 64-bit.  n = 2^23 vertices x 3 walks x length 255 is 6.4 G
    corpus cells (25.7 GB): everything above walk 16.8 M lives beyond 2^32.  The graph is small (an RMAT scale-12 block on
    the LOWEST ids, a copy on the HIGHEST ids - the walks whose keys exceed 2^32 - and edges across), the rest of the
    vertices are isolated, so the 64-bit C oracle (oracle/wharf_oracle.c) finishes in seconds.  Compared: step counts,
    corpus fingerprints, affected ids and positions after a batch on the high block, and walks read back from both ends."""
    import psutil
    from oracle import wharf_oracle as wo
    if psutil.virtual_memory().available < (80 << 30):
        pytest.skip("needs 80 GB of host memory for the oracle's corpus")
    scale, block = 23, 1 << 12
    n = 1 << scale
    off12, adj12 = wo.rmat_graph(12, 16000, 5)
    low = np.stack([np.repeat(np.arange(block, dtype=np.int64), np.diff(np.append(off12, adj12.shape[0]).astype(np.int64))), adj12.astype(np.int64)], 1)
    high = low + (n - block)
    rng = np.random.default_rng(11)
    cross = np.stack([rng.integers(0, block, 3000), rng.integers(n - block, n, 3000)], 1)
    pairs = np.concatenate([low, high, cross, cross[:, ::-1]])
    pairs = np.unique(pairs[pairs[:, 0] != pairs[:, 1]], axis=0)                       # sorted by (src, dst), unique
    off, adj = wo.csr_from_sorted_pairs(n, pairs.astype(np.uint32))
    wpv, length = 3, 255
    assert n * wpv * length > (1 << 32)
    batch = np.stack([rng.integers(n - block, n, 400), rng.integers(n - block, n, 400)], 1).astype(np.uint32)
    batch = np.concatenate([batch, batch[:, ::-1]])
    for det in (True, False):
        model_o, model_e = (wo.DEEPWALK, wharf.DEEPWALK) if det else (wo.NODE2VEC, wharf.NODE2VEC)
        o = wo.Oracle(n, off, adj, walks_per_vertex=wpv, walk_length=length, model=model_o, p=0.5, q=2.0, deterministic=det, seed=3)
        cfg = wharf.Config(walks_per_vertex=wpv, walk_length=length, random_walk_model=model_e, paramP=0.5, paramQ=2.0,
                           deterministic_mode=det, seed=3, search_mode=wharf.SEARCH_SCAN)
        with wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg) as e:
            assert e.generate_initial_random_walks() == o.generate()
            ow = o.walks()
            assert e.corpus_fingerprint() == wo.corpus_fingerprint(ow)
            probe = [0, 5, n - 1, n, 2 * n + n - 7, 3 * n - 1, 2 * n + n - block + 17]   # the last ones: keys beyond 2^32
            assert max(probe) * length > (1 << 32)
            assert np.array_equal(e.walks_gather(probe), ow[probe])
            assert e.walk(3 * n - 1) == "".join(f"{v} " for v in ow[3 * n - 1][ow[3 * n - 1] != wo.NO_VERTEX])
            o.update(batch.copy(), delete=False, remove_dups=True)
            affected = e.insert_edges_batch(batch.copy(), remove_dups=True)
            ids, pos = o.affected()
            assert ids.max() * length > (1 << 32) and np.array_equal(affected, ids)
            assert np.array_equal(e.affected_walks()[1], pos)
            assert e.corpus_fingerprint() == wo.corpus_fingerprint(o.walks())
            assert np.array_equal(e.walks_gather(probe), o.walks()[probe])
        del o, ow
