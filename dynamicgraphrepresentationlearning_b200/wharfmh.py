"""Host-side mirror of the reference's public interface for the hot path.

`WharfMH` follows dygrl::WharfMH (graph/wharfmh.h:21) member for member; `Config` carries the mutable
globals of namespace config (config/globals.h:7-29).  Everything computes in libwharfmh_b200.so through
the C ABI of include/wharfmh_b200.h.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _capi as capi
from ._capi import (DEEPWALK, INIT_BURNIN, INIT_RANDOM, INIT_WEIGHT, NODE2VEC, NO_VERTEX, REF_SENTINEL, SEARCH_AUTO,
                    SEARCH_INDEX, SEARCH_SCAN, WharfError, WmhConfig, check)


@dataclass
class Config:
    """namespace config (config/globals.h:7-29) + placement of this engine."""
    walks_per_vertex: int = 10            # config::walks_per_vertex
    walk_length: int = 80                 # config::walk_length
    random_walk_model: int = NODE2VEC     # config::random_walk_model
    paramP: float = 4.0                   # config::paramP
    paramQ: float = 1.0                   # config::paramQ
    sampler_init_strategy: int = INIT_WEIGHT
    deterministic_mode: bool = True       # config::deterministic_mode
    seed: int = 0                         # keyed-MH seed (reference: Random(time(nullptr)))
    search_mode: int = SEARCH_AUTO
    device: int = 0
    rank: int = 0
    world: int = 1
    generate_mode: int = 0                # GENERATE_AUTO / GENERATE_PER_WALKER / GENERATE_COALESCED

    def to_c(self) -> WmhConfig:
        return WmhConfig(self.walks_per_vertex, self.walk_length, self.random_walk_model, int(self.deterministic_mode),
                         self.paramP, self.paramQ, self.sampler_init_strategy, self.search_mode, self.seed, self.device,
                         self.rank, self.world, self.generate_mode)


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


class WharfMH:
    """dygrl::WharfMH (graph/wharfmh.h:21) on one B200.

    WharfMH(n, m, offsets, edges)   - graph/wharfmh.h:58 (offsets: n entries, rows sorted & unique)
    WharfMH(n)                      - graph/wharfmh.h:26 (n isolated vertices)
    """

    def __init__(self, graph_vertices: int, graph_edges: int = 0, offsets=None, edges=None, config: Config | None = None,
                 device_arrays: bool = False):
        self._lib = capi.load()
        self.config = config or Config()
        self._h = C.c_void_p()
        cfg = self.config.to_c()
        if offsets is None:
            check(self._lib.wmh_create_empty(C.byref(self._h), graph_vertices, C.byref(cfg)))
        elif device_arrays:
            # offsets / edges are integer device addresses (e.g. torch.Tensor.data_ptr())
            check(self._lib.wmh_create_device(C.byref(self._h), graph_vertices, graph_edges, C.c_void_p(int(offsets)),
                                              C.c_void_p(int(edges)), C.byref(cfg)))
        else:
            off = np.ascontiguousarray(offsets, dtype=np.uint64)
            adj = np.ascontiguousarray(edges, dtype=np.uint32)
            if off.shape[0] < graph_vertices or adj.shape[0] < graph_edges:
                raise ValueError("offsets needs n entries and edges m entries")
            check(self._lib.wmh_create(C.byref(self._h), graph_vertices, graph_edges, _ptr(off), _ptr(adj), C.byref(cfg)))

    # -- lifecycle ------------------------------------------------------------------------------------
    def destroy(self) -> None:                                  # graph/wharfmh.h:228
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.wmh_destroy(self._h)
            self._h = C.c_void_p()

    close = destroy

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.destroy()

    def destroy_index(self) -> None:                            # graph/wharfmh.h:237
        check(self._lib.wmh_destroy_index(self._h))

    def set_config(self, **changes) -> None:
        """Assign config:: knobs between calls (experiments/src/throughput-latency.cpp:19-75)."""
        for k, v in changes.items():
            if not hasattr(self.config, k):
                raise AttributeError(k)
            setattr(self.config, k, v)
        cfg = self.config.to_c()
        check(self._lib.wmh_set_config(self._h, C.byref(cfg)))

    # -- graph ----------------------------------------------------------------------------------------
    def number_of_vertices(self) -> int:                        # graph/wharfmh.h:117
        return int(self._lib.wmh_number_of_vertices(self._h))

    def number_of_edges(self) -> int:                           # graph/wharfmh.h:130
        return int(self._lib.wmh_number_of_edges(self._h))

    def flatten_graph(self) -> tuple[np.ndarray, np.ndarray]:   # graph/wharfmh.h:175
        n, m = self.number_of_vertices(), self.number_of_edges()
        row_ptr = np.empty(n + 1, dtype=np.uint64)
        col = np.empty(m, dtype=np.uint32)
        check(self._lib.wmh_flatten_graph(self._h, _ptr(row_ptr), _ptr(col)))
        return row_ptr, col

    def degree(self, vertex: int) -> int:
        d = C.c_uint64()
        check(self._lib.wmh_degree(self._h, vertex, C.byref(d)))
        return int(d.value)

    def neighbors(self, vertex: int) -> np.ndarray:             # graph/snapshot.h:187
        out = np.empty(self.degree(vertex), dtype=np.uint32)
        d = C.c_uint64()
        check(self._lib.wmh_neighbors(self._h, vertex, _ptr(out), out.shape[0], C.byref(d)))
        return out

    # -- walks ----------------------------------------------------------------------------------------
    def generate_initial_random_walks(self) -> int:             # graph/wharfmh.h:250
        check(self._lib.wmh_generate_initial_random_walks(self._h))
        return self.last_steps()

    def generate_initial_random_walks_to_host(self, out: np.ndarray) -> int:
        """generate_initial_random_walks with the corpus streamed into `out` ([local_walks, walk_length] uint32, C order;
        pinned memory overlaps the copies with the walk kernels).  Returns the walker steps."""
        if out.dtype != np.uint32 or not out.flags["C_CONTIGUOUS"] or out.ndim != 2 or out.shape[1] != self.config.walk_length:
            raise ValueError("out must be a C-contiguous uint32 array [local_walks, walk_length]")
        check(self._lib.wmh_generate_initial_random_walks_to_host(self._h, _ptr(out), out.shape[0]))
        return self.last_steps()

    def last_steps(self) -> int:
        return int(self._lib.wmh_last_steps(self._h))

    def number_of_walks(self) -> int:
        return int(self._lib.wmh_number_of_walks(self._h))

    def number_of_local_walks(self) -> int:
        return int(self._lib.wmh_number_of_local_walks(self._h))

    def _batch(self, fn, edges, sorted_, remove_dups, nn, apply_walk_updates) -> np.ndarray:
        pairs = edges if (isinstance(edges, np.ndarray) and edges.dtype == np.uint32 and edges.flags.c_contiguous) \
            else np.ascontiguousarray(edges, dtype=np.uint32)
        pairs = pairs.reshape(-1, 2)
        k = C.c_uint64()
        check(fn(self._h, pairs.shape[0], _ptr(pairs), int(sorted_), int(remove_dups),
                 (1 << 64) - 1 if nn is None else nn, int(apply_walk_updates), C.byref(k)))
        return self.affected_walks()[0]

    def insert_edges_batch(self, edges, sorted: bool = False, remove_dups: bool = False, nn: int | None = None,
                           apply_walk_updates: bool = True, run_seq: bool = False) -> np.ndarray:
        """graph/wharfmh.h:439.  `edges`: (m, 2) uint32 (src, dst); sorted in place when a C-contiguous
        uint32 array is passed and sorted=False.  Returns the affected (global) walk ids, ascending."""
        return self._batch(self._lib.wmh_insert_edges_batch, edges, sorted, remove_dups, nn, apply_walk_updates)

    def delete_edges_batch(self, edges, sorted: bool = False, remove_dups: bool = False, nn: int | None = None,
                           apply_walk_updates: bool = True, run_seq: bool = False) -> np.ndarray:
        """graph/wharfmh.h:588."""
        return self._batch(self._lib.wmh_delete_edges_batch, edges, sorted, remove_dups, nn, apply_walk_updates)

    def insert_edges_batch_device(self, d_edges: int, m: int, remove_dups: bool = False, apply_walk_updates: bool = True) -> int:
        k = C.c_uint64()
        check(self._lib.wmh_insert_edges_batch_device(self._h, m, C.c_void_p(int(d_edges)), int(remove_dups), int(apply_walk_updates), C.byref(k)))
        return int(k.value)

    def delete_edges_batch_device(self, d_edges: int, m: int, remove_dups: bool = False, apply_walk_updates: bool = True) -> int:
        k = C.c_uint64()
        check(self._lib.wmh_delete_edges_batch_device(self._h, m, C.c_void_p(int(d_edges)), int(remove_dups), int(apply_walk_updates), C.byref(k)))
        return int(k.value)

    def number_of_affected_walks(self) -> int:
        k = C.c_uint64()
        check(self._lib.wmh_affected_walks_device(self._h, None, None, C.byref(k)))
        return int(k.value)

    def affected_walks(self) -> tuple[np.ndarray, np.ndarray]:
        """(global walk ids ascending, earliest affected position) of the last batch (types::MapOfChanges)."""
        k = self.number_of_affected_walks()
        ids = np.empty(k, dtype=np.uint64)
        pos = np.empty(k, dtype=np.uint32)
        check(self._lib.wmh_affected_walks(self._h, _ptr(ids), _ptr(pos), k))
        return ids, pos

    def affected_walks_global_device(self, d_out: int, capacity: int) -> int:
        """Global ids (ascending) of the last batch's affected walks into device memory at `d_out` (u64 x capacity)."""
        k = self.number_of_affected_walks()
        check(self._lib.wmh_affected_walks_global_device(self._h, C.c_void_p(int(d_out)), int(capacity)))
        return k

    def sampler_chain(self, cur: int, prev: int, iters: int) -> tuple[int, np.ndarray]:
        """MetropolisHastingsSampler::sample as a chain on the fixed state (cur, prev) (metropolis_hastings_sampler.h:31-46,
        tests/sampler.cpp:86-127): (init value, counts per neighbour of cur in row order)."""
        degree = self.degree(cur)
        counts = np.zeros(max(degree, 1), dtype=np.uint64)
        init = C.c_uint32()
        check(self._lib.wmh_sampler_chain(self._h, cur, prev, iters, C.byref(init), _ptr(counts), counts.shape[0]))
        return int(init.value), counts[:degree]

    def touched_vertices(self) -> np.ndarray:
        k = C.c_uint64()
        check(self._lib.wmh_touched_vertices(self._h, None, 0, C.byref(k)))
        out = np.empty(int(k.value), dtype=np.uint32)
        check(self._lib.wmh_touched_vertices(self._h, _ptr(out), out.shape[0], C.byref(k)))
        return out

    def scan_touched_no_alloc(self, d_vertices: int, count: int, d_min_pos: int) -> None:
        """wmh_scan_touched_no_alloc: device pointers in and out, no allocation inside (static scratch, <= 2^25 vertices)."""
        check(self._lib.wmh_scan_touched_no_alloc(self._h, C.c_void_p(d_vertices), count, C.c_void_p(d_min_pos)))

    def walk(self, walk_id: int) -> str:                        # graph/wharfmh.h:365
        need = C.c_size_t()
        check(self._lib.wmh_walk_string(self._h, walk_id, None, 0, C.byref(need)))
        buf = C.create_string_buffer(need.value)
        check(self._lib.wmh_walk_string(self._h, walk_id, buf, need.value, C.byref(need)))
        return buf.value.decode()

    def walk_vertices(self, walk_id: int) -> np.ndarray:
        out = np.empty(self.config.walk_length, dtype=np.uint32)
        length = C.c_uint32()
        check(self._lib.wmh_walk(self._h, walk_id, _ptr(out), C.byref(length)))
        return out[: length.value]

    def vertex_at_walk(self, walk_id: int, position: int) -> int:   # graph/wharfmh.h:404
        v = C.c_uint32()
        check(self._lib.wmh_vertex_at_walk(self._h, walk_id, position, C.byref(v)))
        return int(v.value)

    def walks(self, first: int = 0, count: int | None = None) -> np.ndarray:
        """Local walks [first, first+count) as rows of walk_length vertices (NO_VERTEX padded)."""
        if count is None:
            count = self.number_of_local_walks() - first
        out = np.empty((count, self.config.walk_length), dtype=np.uint32)
        check(self._lib.wmh_walks_copy(self._h, first, count, _ptr(out)))
        return out

    def walks_gather(self, local_ids) -> np.ndarray:
        ids = np.ascontiguousarray(local_ids, dtype=np.uint32)
        out = np.empty((ids.shape[0], self.config.walk_length), dtype=np.uint32)
        check(self._lib.wmh_walks_gather(self._h, _ptr(ids), ids.shape[0], _ptr(out)))
        return out

    def write_walks(self, path, walk_ids, append: bool = False) -> int:
        """walks.txt (experiments/src/vertex-classification.cpp:167-178): walk(id) + newline for every id (global ids
        this engine holds), gathered from the device in bulk.  Returns the bytes written."""
        import os
        ids = np.ascontiguousarray(walk_ids, dtype=np.uint64)
        written = C.c_uint64()
        check(self._lib.wmh_write_walks(self._h, _ptr(ids), ids.shape[0], os.fspath(path).encode(), int(append), C.byref(written)))
        return int(written.value)

    def local_to_global_walk(self, local: int) -> int:
        return int(self._lib.wmh_local_to_global_walk(self._h, local))

    def global_to_local_walk(self, walk_id: int) -> int:
        return int(self._lib.wmh_global_to_local_walk(self._h, walk_id))

    def local_walk_ids(self) -> np.ndarray:
        """Global ids of the local walks in local order (ascending): the sharding rule of include/wharfmh_b200.h."""
        count = int(self._lib.wmh_local_walk_ids(self._h, None, 0))
        out = np.empty(count, dtype=np.uint64)
        if count:
            self._lib.wmh_local_walk_ids(self._h, out.ctypes.data_as(C.c_void_p), count)
        return out

    def walk_owner(self, walk_id: int) -> int:
        """Rank that owns a global walk id."""
        return int(self._lib.wmh_walk_owner(self._h, int(walk_id)))

    def corpus_fingerprint(self) -> int:
        h = C.c_uint64()
        check(self._lib.wmh_corpus_fingerprint(self._h, C.byref(h)))
        return int(h.value)

    # -- inverted index / paired corpus ---------------------------------------------------------------
    def build_index(self) -> None:
        check(self._lib.wmh_build_index(self._h))

    def find_next(self, vertex: int, walk_id: int, position: int) -> tuple[int, bool]:
        """InvertedIndex::find_next (walks/inverted_index.h:74-92) on `vertex`."""
        nxt, found = C.c_uint32(), C.c_int()
        check(self._lib.wmh_index_find_next(self._h, vertex, walk_id, position, C.byref(nxt), C.byref(found)))
        return int(nxt.value), bool(found.value)

    def index_sizes(self) -> np.ndarray:
        """Entries of every vertex's inverted index (all zero while no corpus exists)."""
        out = np.empty(self.number_of_vertices(), dtype=np.uint64)
        check(self._lib.wmh_index_sizes(self._h, _ptr(out)))
        return out

    def index_size(self, vertex: int) -> int:
        k = C.c_uint64()
        check(self._lib.wmh_index_size(self._h, vertex, C.byref(k)))
        return int(k.value)

    def compressed_walks(self) -> tuple[np.ndarray, np.ndarray]:
        """CompressedWalks (walks/compressed_walks.h): (ptr[n+1], codes) with codes = pair(walk_id*L+pos, next)."""
        total = C.c_uint64()
        ptr = np.empty(self.number_of_vertices() + 1, dtype=np.uint64)
        check(self._lib.wmh_compressed_walks(self._h, _ptr(ptr), None, 0, C.byref(total)))
        codes = np.empty(int(total.value), dtype=np.uint64)
        check(self._lib.wmh_compressed_walks(self._h, _ptr(ptr), _ptr(codes), codes.shape[0], C.byref(total)))
        return ptr, codes

    def compressed_find_next(self, vertex: int, walk_id: int, position: int) -> tuple[int, bool]:
        nxt, found = C.c_uint32(), C.c_int()
        check(self._lib.wmh_compressed_find_next(self._h, vertex, walk_id, position, C.byref(nxt), C.byref(found)))
        return int(nxt.value), bool(found.value)

    # -- reporting ------------------------------------------------------------------------------------
    def timer(self, which: int) -> float:                       # config/config.h:10-14
        return float(self._lib.wmh_timer_get(self._h, which))

    def reset_timers(self) -> None:
        for t in range(5):
            self._lib.wmh_timer_reset(self._h, t)

    def memory_footprint(self) -> dict:                         # graph/wharfmh.h:928
        mem = capi.WmhMemory()
        check(self._lib.wmh_memory_footprint(self._h, C.byref(mem)))
        return {n: int(getattr(mem, n)) for n, _ in mem._fields_}

    def kernel_launches(self) -> int:
        return int(self._lib.wmh_kernel_launches(self._h))

    def last_walk_kernel_ms(self) -> float:
        return float(self._lib.wmh_last_walk_kernel_ms(self._h))

    def last_index_ms(self) -> float:
        return float(self._lib.wmh_last_index_ms(self._h))

    def last_search_mode(self) -> int:
        """SEARCH_INDEX / SEARCH_SCAN as taken by the last batch (0: none)."""
        return int(self._lib.wmh_last_search_mode(self._h))

    def neighbour_filter(self) -> int:
        """0: no aux words yet, 1: per-row blocks, 2: triangle blocks (wmh_neighbour_filter)."""
        return int(self._lib.wmh_neighbour_filter(self._h))

    def filter_check(self) -> int:
        """Edge slots whose membership filter lacks a bit it must carry (wmh_filter_check): 0 on a sound engine."""
        missing = C.c_uint64()
        check(self._lib.wmh_filter_check(self._h, C.byref(missing)))
        return int(missing.value)

    def device_walks_ptr(self) -> int:
        return int(self._lib.wmh_device_walks(self._h) or 0)

    def synchronize(self) -> None:
        check(self._lib.wmh_synchronize(self._h))

    # -- peer memory between the ranks of one box (include/wharfmh_b200.h, wmh_peer_*) ------------------
    def peer_bytes_needed(self) -> int:
        return int(self._lib.wmh_peer_bytes_needed(self._h))

    def peer_create(self, nbytes: int) -> bytes:
        """Allocates this rank's exchange buffer; returns its 64-byte CUDA IPC handle."""
        handle = (C.c_ubyte * 64)()
        check(self._lib.wmh_peer_create(self._h, int(nbytes), handle))
        return bytes(handle)

    def peer_attach(self, handles: list[bytes]) -> None:
        """Maps the exchange buffers of all ranks (handles in rank order, 64 bytes each)."""
        blob = b"".join(handles)
        if len(blob) != 64 * len(handles):
            raise ValueError("every handle must be 64 bytes")
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        check(self._lib.wmh_peer_attach(self._h, buf))

    def peer_detach(self) -> None:
        check(self._lib.wmh_peer_detach(self._h))

    def peer_exchanges(self) -> int:
        return int(self._lib.wmh_peer_exchanges(self._h))


def trim_memory(device: int = 0) -> None:
    """Returns the blocks cached by destroyed engines / finished batches to the driver (wmh_trim_memory)."""
    check(capi.load().wmh_trim_memory(device))


# ---- pairings (walks/pairings.h) over torch / raw device pointers ------------------------------------

def generate_batch_of_edges(edges_number: int, vertices_number: int, batch_seed: int, self_loops: bool = False, directed: bool = True,
                            a: float = 0.5, b: float = 0.2, c: float = 0.1, device: int = 0) -> np.ndarray:
    """utility::generate_batch_of_edges (utils/utility.h:55-146), produced on the device: (k, 2) uint32 pairs sorted by
    (src, dst), duplicates (and, unless self_loops, loops) removed - the same bytes the reference's generator returns."""
    slots = edges_number * (1 if directed else 2)
    out = np.empty((slots, 2), dtype=np.uint32)
    k = C.c_uint64()
    check(capi.load().wmh_generate_batch_of_edges(device, edges_number, vertices_number, batch_seed, int(self_loops), int(directed), a, b, c,
                                                  _ptr(out), slots, C.byref(k)))
    return out[: int(k.value)].copy()


def random_stream(seed: int, count: int) -> np.ndarray:
    """The first `count` lrand() of utility::Random(seed) (utils/utility.h:152-206): row `seed` of the draw table."""
    out = np.empty(count, dtype=np.uint64)
    capi.load().wmh_random_stream(seed, count, _ptr(out))
    return out


def unit_double(draw: int) -> float:
    """utility::Random::drand applied to one 64-bit draw (utils/utility.h:208-218)."""
    return float(capi.load().wmh_unit_double(draw))


def keyed_draw(seed: int, domain: int, a: int, b: int, c: int) -> int:
    """The counter-based draw of keyed MH mode (include/wharfmh_b200.h, wmh_keyed_draw)."""
    return int(capi.load().wmh_keyed_draw(seed, domain, a, b, c))


def partition_owner(n: int, walks_per_vertex: int, world: int, walk_id: int) -> int:
    """Rank that owns a global walk id under the sharding rule of include/wharfmh_b200.h (host arithmetic, no device)."""
    return int(capi.load().wmh_partition_owner(int(n), int(walks_per_vertex), int(world), int(walk_id)))


def partition_local_walk_ids(n: int, walks_per_vertex: int, rank: int, world: int) -> np.ndarray:
    """Global ids of the walks `rank` of `world` owns, ascending (host arithmetic, no device)."""
    lib = capi.load()
    count = int(lib.wmh_partition_local_walk_ids(int(n), int(walks_per_vertex), int(rank), int(world), None, 0))
    out = np.empty(count, dtype=np.uint64)
    if count:
        lib.wmh_partition_local_walk_ids(int(n), int(walks_per_vertex), int(rank), int(world), out.ctypes.data_as(C.c_void_p), count)
    return out


def szudzik_pair(x: int, y: int) -> int:
    return int(capi.load().wmh_szudzik_pair(x, y))


def szudzik_unpair(z: int) -> tuple[int, int]:
    x, y = C.c_uint64(), C.c_uint64()
    capi.load().wmh_szudzik_unpair(z, C.byref(x), C.byref(y))
    return int(x.value), int(y.value)
