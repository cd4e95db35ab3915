"""Multi-GPU sharding of the walk corpus: one process per GPU, torch.distributed (NCCL over NVLink) for the
two exchanges the path has.

The reference is single-process; this is the B200 scale-out of the same API.  Walks shard by
replica first (walks_per_vertex // world whole replicas per rank) and the remaining replicas by start vertex (one
vertex of every block of `world` consecutive ids per rank, rotated from block to block: wmh_walk_owner /
wmh_local_walk_ids), every rank holds a full replica of the graph, and no collective runs while walking.  Per edge batch:
  C1  broadcast of the batch from the source rank (<= 16 MB for 2 M edges), so every replica merges the
      same edges;
  C2  all-gather of the affected walk ids (counts first, then a padded all-gather) so that every rank can
      return the global `affected_walks` sequence of graph/wharfmh.h:439 / :588.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from .wharfmh import Config, WharfMH


def connect_peers(engine, group=None, slack: float = 1.25) -> bool:
    """Gives the ranks of one box a window into each other's memory (include/wharfmh_b200.h, wmh_peer_*): every rank
    allocates an exchange buffer, the 64-byte CUDA IPC handles travel through torch.distributed (host plumbing), and
    every rank maps the others'.  From then on work that is a pure function of the replicated graph (node2vec sampler
    init values) is computed once per box and stored into the peers over NVLink by the kernel that computes it.
    Returns False when the engine's configuration has nothing to exchange.  Collective: call it on every rank."""
    need = engine.peer_bytes_needed()
    sizes = [None] * dist.get_world_size(group)
    dist.all_gather_object(sizes, need, group=group)
    if max(sizes) == 0:
        return False
    nbytes = int(max(sizes) * slack) + (1 << 20)        # room for the graph to grow by edge batches
    handle = engine.peer_create(nbytes)
    handles = [None] * dist.get_world_size(group)
    dist.all_gather_object(handles, handle, group=group)   # also orders every create before any attach
    engine.peer_attach(handles)
    dist.barrier(group=group)
    return True


class ShardedWharfMH:
    """dygrl::WharfMH over `world` ranks.  `engine_factory(config) -> engine` builds the per-rank engine (default:
    the CUDA engine over the given host CSR); tests inject a stand-in to exercise the exchange logic on gloo."""

    def __init__(self, graph_vertices: int, graph_edges: int, offsets, edges, config: Config | None = None, group=None,
                 engine_factory=None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised (one process per GPU)")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        cfg = config or Config()
        cfg.rank, cfg.world = self.rank, self.world
        self.config = cfg
        self._on_gpu = dist.get_backend(group) == "nccl"
        self.device = torch.device(f"cuda:{cfg.device}") if self._on_gpu else torch.device("cpu")
        if engine_factory is None:
            self.engine = WharfMH(graph_vertices, graph_edges, offsets, edges, config=cfg)
            if self._on_gpu and 1 < self.world <= 8:
                connect_peers(self.engine, group)
        else:
            self.engine = engine_factory(cfg)
        self.last_timing: dict = {}

    @classmethod
    def from_engine(cls, engine: WharfMH, group=None) -> "ShardedWharfMH":
        """Wraps the engine this rank already built (its config carries rank/world): the exchanges run on the same object
        that generated the corpus."""
        self = cls.__new__(cls)
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if engine.config.rank != self.rank or engine.config.world != self.world:
            raise ValueError("the engine was created for another rank / world")
        self.config = engine.config
        self._on_gpu = dist.get_backend(group) == "nccl"
        self.device = torch.device(f"cuda:{engine.config.device}") if self._on_gpu else torch.device("cpu")
        self.engine = engine
        self.last_timing = {}
        return self

    # -- helpers --------------------------------------------------------------------------------------
    def _all_reduce_sum(self, value: int) -> int:
        t = torch.tensor([value], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return int(t.item())

    def broadcast_batch(self, edges, src: int = 0) -> torch.Tensor:
        """C1: returns the (m, 2) int32 batch on every rank (device tensor under NCCL)."""
        count = torch.zeros(1, dtype=torch.int64, device=self.device)
        if self.rank == src:
            pairs = np.ascontiguousarray(edges, dtype=np.uint32).reshape(-1, 2)
            count[0] = pairs.shape[0]
        dist.broadcast(count, src=src, group=self.group)
        m = int(count.item())
        if self.rank == src:
            t = torch.from_numpy(pairs.view(np.int32)).to(self.device)
        else:
            t = torch.empty((m, 2), dtype=torch.int32, device=self.device)
        if m:
            dist.broadcast(t, src=src, group=self.group)
        return t

    def all_gather_ids(self, local_ids: np.ndarray) -> np.ndarray:
        """C2: variable-length all-gather of u64 ids: counts, then one padded all-gather; returns them sorted."""
        counts = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        mine = torch.tensor([local_ids.shape[0]], dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(counts, mine, group=self.group)
        width = int(counts.max().item())
        if width == 0:
            return np.empty(0, dtype=np.uint64)
        padded = torch.zeros(width, dtype=torch.int64, device=self.device)
        padded[: local_ids.shape[0]] = torch.from_numpy(local_ids.astype(np.int64)).to(self.device)
        gathered = torch.empty(self.world * width, dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(gathered, padded, group=self.group)
        parts = [gathered[r * width: r * width + int(counts[r].item())] for r in range(self.world)]
        return np.sort(torch.cat(parts).cpu().numpy().astype(np.uint64))

    def warm_collectives(self, ids_per_rank: int = 4) -> None:
        """One untimed round of the collectives an update uses (broadcast, all-gather of `ids_per_rank` ids, device sort): NCCL
        opens its channels for each collective kind on first use and sizes its buffers on the first large message (71 ms for
        the first all-gather at N = 2, 20 ms for the first 80 MB one at N = 8), which is set-up, not batch latency."""
        count = torch.zeros(1, dtype=torch.int64, device=self.device)
        dist.broadcast(count, src=0, group=self.group)
        pairs = torch.zeros((16, 2), dtype=torch.int32, device=self.device)
        dist.broadcast(pairs, src=0, group=self.group)
        ids_per_rank = max(1, int(ids_per_rank))
        ids = torch.arange(ids_per_rank, dtype=torch.int64, device=self.device) + ids_per_rank * self.rank
        if self._on_gpu:
            self.all_gather_ids_device(ids)
            torch.cuda.synchronize(self.device)
        else:
            self.all_gather_ids(ids.numpy().astype(np.uint64))

    # -- dygrl::WharfMH surface -----------------------------------------------------------------------
    def number_of_vertices(self) -> int:
        return self.engine.number_of_vertices()

    def number_of_edges(self) -> int:
        return self.engine.number_of_edges()

    def generate_initial_random_walks(self) -> int:
        """graph/wharfmh.h:250 on every shard; returns the global number of walker steps."""
        return self._all_reduce_sum(self.engine.generate_initial_random_walks())

    def all_gather_ids_device(self, local: torch.Tensor) -> torch.Tensor:
        """C2 without leaving the device: `local` = this rank's affected global ids (int64 device tensor); returns the
        ascending ids of all ranks as a device tensor (counts, one padded NCCL all-gather, device sort)."""
        counts = torch.zeros(self.world, dtype=torch.int64, device=self.device)
        mine = torch.tensor([local.shape[0]], dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(counts, mine, group=self.group)
        host_counts = counts.tolist()
        width = max(host_counts)
        if width == 0:
            return torch.empty(0, dtype=torch.int64, device=self.device)
        padded = torch.empty(width, dtype=torch.int64, device=self.device)
        padded[: local.shape[0]] = local
        gathered = torch.empty(self.world * width, dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(gathered, padded, group=self.group)
        parts = [gathered[r * width: r * width + host_counts[r]] for r in range(self.world)]
        return torch.sort(torch.cat(parts)).values

    def _update(self, edges, delete: bool, remove_dups: bool, apply_walk_updates: bool, src: int, return_device: bool = False):
        import time
        if not self._on_gpu:
            batch = self.broadcast_batch(edges, src)
            fn = self.engine.delete_edges_batch if delete else self.engine.insert_edges_batch
            local = fn(batch.numpy().view(np.uint32).copy(), sorted=False, remove_dups=remove_dups, apply_walk_updates=apply_walk_updates)
            return self.all_gather_ids(np.asarray(local, dtype=np.uint64))
        t0 = time.perf_counter()
        batch = self.broadcast_batch(edges, src)                  # C1
        torch.cuda.current_stream().synchronize()                  # the engine runs on its own stream
        t1 = time.perf_counter()
        fn = self.engine.delete_edges_batch_device if delete else self.engine.insert_edges_batch_device
        k = fn(batch.data_ptr(), int(batch.shape[0]), remove_dups=remove_dups, apply_walk_updates=apply_walk_updates)
        t2 = time.perf_counter()
        local = torch.empty(max(k, 1), dtype=torch.int64, device=self.device)
        self.engine.affected_walks_global_device(local.data_ptr(), local.shape[0])
        ids = self.all_gather_ids_device(local[:k])                # C2
        torch.cuda.current_stream().synchronize()
        t3 = time.perf_counter()
        self.last_timing = {"bcast_ms": 1e3 * (t1 - t0), "update_ms": 1e3 * (t2 - t1), "allgather_ms": 1e3 * (t3 - t2),
                            "local_affected": int(k), "global_affected": int(ids.shape[0])}
        if return_device:
            return ids
        return ids.cpu().numpy().astype(np.uint64)

    def insert_edges_batch(self, edges=None, remove_dups: bool = False, apply_walk_updates: bool = True, src: int = 0,
                           return_device: bool = False):
        """graph/wharfmh.h:439: `edges` is read on rank `src` only; returns the global affected walk ids (ascending) on
        every rank - a numpy array, or the device tensor the all-gather produced when return_device.  Under NCCL
        `last_timing` holds the milliseconds of the broadcast, the local update and the all-gather."""
        return self._update(edges, False, remove_dups, apply_walk_updates, src, return_device)

    def delete_edges_batch(self, edges=None, remove_dups: bool = False, apply_walk_updates: bool = True, src: int = 0,
                           return_device: bool = False):
        """graph/wharfmh.h:588."""
        return self._update(edges, True, remove_dups, apply_walk_updates, src, return_device)

    def walk(self, walk_id: int) -> str:
        """graph/wharfmh.h:365 for any global walk id: the owner broadcasts the text."""
        owner = self.engine.walk_owner(walk_id)
        box = [self.engine.walk(walk_id) if self.rank == owner else None]
        dist.broadcast_object_list(box, src=owner, group=self.group)
        return box[0]

    def write_walks(self, path, walk_ids, append: bool = False) -> int:
        """The walks.txt hand-off (experiments/src/vertex-classification.cpp:167-178) of a sharded corpus: `walk_ids` are
        global ids (what insert/delete_edges_batch returned, the same on every rank); every rank writes the walks it
        owns, one rank after the other, into the same file - the ranks of a box share the file system, and no walk
        crosses a link.  Lines come out grouped by rank, ascending id inside a group (the reference's order is its hash
        map's).  Returns the bytes written by all ranks.  Collective."""
        ids = np.asarray(walk_ids, dtype=np.uint64)
        mine = ids[np.isin(ids, self.engine.local_walk_ids())]
        written = 0
        for turn in range(self.world):
            if turn == self.rank:
                written = self.engine.write_walks(path, np.sort(mine), append=append or turn > 0)
            dist.barrier(group=self.group)
        return self._all_reduce_sum(written)

    def local_walks(self) -> tuple[np.ndarray, np.ndarray]:
        """(global ids, rows) of the walks this rank owns."""
        return self.engine.local_walk_ids(), self.engine.walks()

    def destroy(self) -> None:
        self.engine.destroy()
