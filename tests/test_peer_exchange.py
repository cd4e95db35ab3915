"""Two ranks on two GPUs of one box: peer-memory exchange (include/wharfmh_b200.h, wmh_peer_*).

The exchange only moves values that are pure functions of the replicated graph, so the corpus must be bit-identical
to the keyed CPU restatement (and to a stand-alone engine) whether peers are attached or not.  Needs >= 2 CUDA
devices; skipped on a single-GPU box (run with `gpurun --gpus 2 -- python -m pytest tests -m gpu`)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(f"cuda:{rank}"))
    try:
        import dynamicgraphrepresentationlearning_b200 as wharf
        from dynamicgraphrepresentationlearning_b200.distributed import ShardedWharfMH, connect_peers
        from oracle import wharf_oracle as wo
        report = []
        for scale, edges, init in ((9, 4000, wo.INIT_WEIGHT), (12, 40000, wo.INIT_WEIGHT), (9, 4000, wo.INIT_RANDOM), (8, 1500, wo.INIT_BURNIN)):
            n = 1 << scale
            off, adj = wo.rmat_graph(scale, edges, seed=11)
            kw = dict(walks_per_vertex=10, walk_length=40)      # large enough a corpus for the exchange to be worth it (engine rule)
            o = wo.Oracle(n, off, adj, model=wo.NODE2VEC, p=0.5, q=2.0, deterministic=False, init_strategy=init, seed=77, **kw)
            cfg = wharf.Config(random_walk_model=wharf.NODE2VEC, paramP=0.5, paramQ=2.0, deterministic_mode=False,
                               sampler_init_strategy=init, seed=77, device=rank, rank=rank, world=world, **kw)
            e = wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=cfg)
            mine = e.local_walk_ids().astype(np.int64)
            # stand-alone first (no peers): lazy sampler init
            e.generate_initial_random_walks()
            alone = e.walks().copy()
            assert e.peer_exchanges() == 0
            assert connect_peers(e)
            o.generate()
            for _ in range(2):                                   # twice: the buffers are reused
                e.generate_initial_random_walks()
                assert np.array_equal(e.walks(), o.walks()[mine]), "corpus with peer exchange differs from the oracle"
            assert np.array_equal(alone, o.walks()[mine])
            assert e.peer_exchanges() == 2
            # batches between generates: re-walks read the shared caches, later generates share again
            for s, delete in ((0, False), (1, False), (0, True)):
                b = wo.generate_batch(25, n, s, False, False)
                o.update(b, delete=delete, remove_dups=True)
                (e.delete_edges_batch if delete else e.insert_edges_batch)(b.copy(), remove_dups=True)
                assert np.array_equal(e.walks(), o.walks()[mine]), "corpus differs after a batch"
            shared = e.peer_exchanges()                          # 2, plus one per batch that rewrote a quarter of the slots
            assert 2 <= shared <= 5
            e.destroy_index()
            o.generate()
            e.generate_initial_random_walks()
            assert np.array_equal(e.walks(), o.walks()[mine])
            assert e.peer_exchanges() == shared + 1
            e.peer_detach()
            e.generate_initial_random_walks()
            assert np.array_equal(e.walks(), o.walks()[mine])
            e.destroy()
            o.close()
            report.append(scale)
        # the wrapper connects its ranks itself
        n = 1 << 9
        off, adj = wo.rmat_graph(9, 4000, seed=5)
        cfg = wharf.Config(walks_per_vertex=10, walk_length=30, random_walk_model=wharf.NODE2VEC, paramP=0.7, paramQ=0.2,
                           deterministic_mode=False, seed=5, device=rank)
        sharded = ShardedWharfMH(n, int(adj.shape[0]), off, adj, config=cfg)
        o = wo.Oracle(n, off, adj, walks_per_vertex=10, walk_length=30, model=wo.NODE2VEC, p=0.7, q=0.2, deterministic=False, seed=5)
        assert sharded.generate_initial_random_walks() == o.generate()
        gids, rows = sharded.local_walks()
        assert np.array_equal(rows, o.walks()[gids.astype(np.int64)])
        assert sharded.engine.peer_exchanges() == 1
        b = wo.generate_batch(20, n, 3, False, False)
        o.update(b, delete=False, remove_dups=True)
        ids = sharded.insert_edges_batch(b if rank == 0 else None, remove_dups=True)
        assert np.array_equal(ids, o.affected()[0])
        gids, rows = sharded.local_walks()
        assert np.array_equal(rows, o.walks()[gids.astype(np.int64)])
        sharded.destroy()
        # a corpus too small to touch most sampler states keeps the lazy first-touch init (no exchange, same corpus)
        cfg = wharf.Config(walks_per_vertex=1, walk_length=4, random_walk_model=wharf.NODE2VEC, paramP=0.7, paramQ=0.2,
                           deterministic_mode=False, seed=5, device=rank)
        tiny = ShardedWharfMH(n, int(adj.shape[0]), off, adj, config=cfg)
        o = wo.Oracle(n, off, adj, walks_per_vertex=1, walk_length=4, model=wo.NODE2VEC, p=0.7, q=0.2, deterministic=False, seed=5)
        assert tiny.generate_initial_random_walks() == o.generate()
        gids, rows = tiny.local_walks()
        assert np.array_equal(rows, o.walks()[gids.astype(np.int64)])
        assert tiny.engine.peer_exchanges() == 0
        tiny.destroy()
        (Path(out_dir) / f"ok{rank}").write_text("True")
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_peer_exchange_two_gpus(tmp_path, wo):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").read_text() == "True" and (tmp_path / "ok1").read_text() == "True"
