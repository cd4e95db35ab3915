"""wmh_scan_touched_no_alloc: the affected-walk corpus scan with its scratch (the touched-vertex bitmap) in a static device
array - an entry point that allocates nothing, for callers that cannot call cudaMalloc.  No counterpart in the reference.
Its output is compared, bit for bit, with what the existing entry points
report for the same batch (wmh_insert/delete_edges_batch -> wmh_affected_walks, wmh_touched_vertices), on calls in a row:
the static array keeps the previous call's contents and must not leak them.

Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NO_POS = 0xFFFFFFFF


def _scan(torch, engine, vertices):
    """min_pos of every local walk through the new entry point (device buffers owned by the caller)."""
    d_vertices = torch.from_numpy(np.ascontiguousarray(vertices, dtype=np.uint32).view(np.int32)).cuda()
    d_min_pos = torch.full((engine.number_of_local_walks(),), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    engine.scan_touched_no_alloc(d_vertices.data_ptr() if d_vertices.numel() else 0, int(d_vertices.numel()), d_min_pos.data_ptr())
    return d_min_pos.cpu().numpy().view(np.uint32)


def test_no_alloc_scan_equals_the_batch_search(wharf, wo):
    import torch
    rng = np.random.default_rng(3)
    n = 1 << 10
    off, adj = wo.rmat_graph(10, 8000, seed=6)
    for search in (wharf.SEARCH_SCAN, wharf.SEARCH_INDEX):
        cfg = wharf.Config(walks_per_vertex=4, walk_length=23, search_mode=search)
        with wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg) as e:
            e.generate_initial_random_walks()
            corpus = e.walks()
            for call, (size, delete) in enumerate(((60, False), (5, False), (200, True), (1, False))):
                batch = rng.integers(0, n, (size, 2)).astype(np.uint32)
                # apply_walk_updates=False: the graph changes, the corpus does not - the search result describes the corpus we scan
                fn = e.delete_edges_batch if delete else e.insert_edges_batch
                fn(batch.copy(), remove_dups=True, apply_walk_updates=False)
                ids, positions = e.affected_walks()
                touched = e.touched_vertices()
                got = _scan(torch, e, touched)
                want = np.full(corpus.shape[0], NO_POS, dtype=np.uint32)
                want[ids.astype(np.int64)] = positions
                assert np.array_equal(got, want), f"call {call}: static-scratch scan differs from wmh_affected_walks"
                # and against the corpus itself
                hit = np.isin(corpus, touched) & (corpus != wharf.NO_VERTEX)
                first = np.where(hit.any(axis=1), hit.argmax(axis=1), NO_POS).astype(np.uint32)
                assert np.array_equal(got, first), f"call {call}"
            # nothing marked: nothing of the previous calls may be left in the array; ids beyond the graph are ignored
            assert np.all(_scan(torch, e, np.empty(0, dtype=np.uint32)) == NO_POS)
            assert np.all(_scan(torch, e, np.array([n, n + 77, 0xFFFFFFFE], dtype=np.uint32)) == NO_POS)
            assert np.array_equal(e.walks(), corpus)


def test_no_alloc_scan_refuses_graphs_beyond_its_scratch(wharf):
    import torch
    n = (1 << 25) + 1                                            # one vertex more than the static bitmap holds
    cfg = wharf.Config(walks_per_vertex=1, walk_length=1)
    with wharf.WharfMH(n, config=cfg) as e:
        e.generate_initial_random_walks()
        d_out = torch.empty(n, dtype=torch.int32, device="cuda")
        with pytest.raises(wharf.WharfError) as err:
            e.scan_touched_no_alloc(0, 0, d_out.data_ptr())
        assert "2^25" in str(err.value)
    with wharf.WharfMH(64, config=cfg) as e:                     # and a corpus that does not exist yet
        d_out = torch.empty(64, dtype=torch.int32, device="cuda")
        with pytest.raises(wharf.WharfError):
            e.scan_touched_no_alloc(0, 0, d_out.data_ptr())
