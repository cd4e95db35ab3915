"""CPU tests (gloo, world_size 2) of the multi-rank host logic: batch broadcast (C1), variable-length all-gather
of affected walk ids (C2) and the walk sharding rule.  The per-rank engine is a stand-in backed by the CPU
oracle (tests may use the oracle; the product never does), sharded exactly like the CUDA engine: the host-side
partition rule of the library itself (wmh_partition_owner / wmh_partition_local_walk_ids)."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]


class OracleShard:
    """Stand-in engine: full oracle per rank, reporting only the walks this rank owns."""

    def __init__(self, wo, n, off, adj, cfg):
        self.o = wo.Oracle(n, off, adj, walks_per_vertex=cfg.walks_per_vertex, walk_length=cfg.walk_length)
        self.cfg, self.n = cfg, n

    def _mine(self, ids):
        return ids[np.isin(ids, self.local_walk_ids())]

    def number_of_vertices(self):
        return self.n

    def number_of_edges(self):
        return self.o.m

    def generate_initial_random_walks(self):
        self.o.generate()
        w = self.o.walks()
        ids = self.local_walk_ids().astype(np.int64)
        return int(((w[ids] != 0xFFFFFFFF).sum(1)[(np.diff(self.o.csr()[0].astype(np.int64))[ids % self.n] > 0)]).sum())

    def _update(self, edges, delete, remove_dups, apply_walk_updates):
        self.o.update(edges, delete=delete, remove_dups=remove_dups, apply_walk_updates=apply_walk_updates)
        return self._mine(self.o.affected()[0])

    def insert_edges_batch(self, edges, sorted=False, remove_dups=False, apply_walk_updates=True):
        return self._update(edges, False, remove_dups, apply_walk_updates)

    def delete_edges_batch(self, edges, sorted=False, remove_dups=False, apply_walk_updates=True):
        return self._update(edges, True, remove_dups, apply_walk_updates)

    def walk_owner(self, walk_id):
        from dynamicgraphrepresentationlearning_b200 import partition_owner
        return partition_owner(self.n, self.cfg.walks_per_vertex, self.cfg.world, walk_id)

    def local_walk_ids(self):
        from dynamicgraphrepresentationlearning_b200 import partition_local_walk_ids
        return partition_local_walk_ids(self.n, self.cfg.walks_per_vertex, self.cfg.rank, self.cfg.world)

    def walks(self):
        return self.o.walks()[self.local_walk_ids().astype(np.int64)]

    def walk(self, walk_id):
        row = self.o.walks()[walk_id]
        return "".join(f"{v} " for v in row[row != 0xFFFFFFFF])

    def write_walks(self, path, walk_ids, append=False):
        from dynamicgraphrepresentationlearning_b200 import graph_io
        ids = np.asarray(walk_ids, dtype=np.int64)
        assert np.isin(ids, self.local_walk_ids()).all()               # the engine refuses walks it does not hold
        text = graph_io.format_walks(self.o.walks()[ids])
        with open(path, "ab" if append else "wb") as f:
            f.write(text)
        return len(text)

    def destroy(self):
        self.o.close()


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dynamicgraphrepresentationlearning_b200.distributed import ShardedWharfMH
        from dynamicgraphrepresentationlearning_b200.wharfmh import Config
        from oracle import wharf_oracle as wo
        n = 256
        off, adj = wo.rmat_graph(8, 1500)
        cfg = Config(walks_per_vertex=3, walk_length=9)
        sharded = ShardedWharfMH(n, adj.shape[0], off, adj, config=cfg, engine_factory=lambda c: OracleShard(wo, n, off, adj, c))
        single = wo.Oracle(n, off, adj, walks_per_vertex=3, walk_length=9)
        assert sharded.generate_initial_random_walks() == single.generate()
        sharded.warm_collectives()                                # an untimed round of the update's collectives: no state change
        ok = True
        for seed, delete in ((0, False), (1, False), (0, True)):
            batch = wo.generate_batch(20, n, seed, False, False)
            single.update(batch, delete=delete, remove_dups=True)
            fn = sharded.delete_edges_batch if delete else sharded.insert_edges_batch
            ids = fn(batch if rank == 0 else None, remove_dups=True)          # only the source rank holds the batch
            ok &= np.array_equal(ids, single.affected()[0])
            ok &= sharded.number_of_edges() == single.m
            gids, rows = sharded.local_walks()
            ok &= np.array_equal(rows, single.walks()[gids.astype(np.int64)])
            ok &= all(sharded.engine.walk_owner(int(g)) == rank for g in gids[::7])
        # walks.txt of the last batch's affected walks: one file, every rank writes its own share in turn
        from dynamicgraphrepresentationlearning_b200 import graph_io, partition_owner
        path = Path(out_dir) / "walks.txt"
        if rank == 0:
            path.write_text("stale\n")
        dist.barrier()
        total = sharded.write_walks(path, ids)
        order = sorted((int(w) for w in ids), key=lambda w: (partition_owner(n, 3, world, w), w))
        want = graph_io.format_walks(single.walks()[np.array(order, dtype=np.int64)])
        ok &= total == len(want) and path.read_bytes() == want
        dist.barrier()                                                   # every rank has read the file before it grows
        ok &= sharded.write_walks(path, ids[:5], append=True) > 0
        ok &= path.read_bytes().startswith(want) and path.read_bytes().count(b"\n") == len(order) + 5
        dist.barrier()
        ok &= sharded.walk(5) == sharded.walk(5) and sharded.walk(258).startswith("2 ")
        empty = sharded.insert_edges_batch(np.empty((0, 2), dtype=np.uint32) if rank == 0 else None)
        ok &= empty.shape[0] == 0
        (Path(out_dir) / f"ok{rank}").write_text(str(bool(ok)))
    finally:
        dist.destroy_process_group()


def test_sharded_host_logic_gloo(tmp_path):
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").read_text() == "True" and (tmp_path / "ok1").read_text() == "True"
