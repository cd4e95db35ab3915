"""ctypes binding of include/wharfmh_b200.h (libwharfmh_b200.so).

The library is the product: if it is missing or cannot be loaded this module raises - there is no
Python or CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "lib" / "libwharfmh_b200.so"

NO_VERTEX = 0xFFFFFFFF
REF_SENTINEL = 0xFFFFFFFE

OK, ERR_INVALID, ERR_CUDA, ERR_OOM, ERR_STATE, ERR_DOMAIN, ERR_RANGE = range(7)
DEEPWALK, NODE2VEC = 0, 1
INIT_RANDOM, INIT_BURNIN, INIT_WEIGHT = 0, 1, 2
SEARCH_AUTO, SEARCH_INDEX, SEARCH_SCAN = 0, 1, 2
GENERATE_AUTO, GENERATE_PER_WALKER, GENERATE_COALESCED = 0, 1, 2
PEER_HANDLE_BYTES, PEER_MAX_WORLD = 64, 8
TIMER_GRAPH_UPDATE_INSERT, TIMER_WALK_UPDATE_INSERT, TIMER_GRAPH_UPDATE_DELETE, TIMER_WALK_UPDATE_DELETE, TIMER_GENERATE = range(5)


class WmhConfig(C.Structure):
    """wmh_config: the knobs of namespace config (config/globals.h:7-29) plus placement."""
    _fields_ = [
        ("walks_per_vertex", C.c_uint32),
        ("walk_length", C.c_uint32),
        ("model", C.c_uint32),
        ("deterministic", C.c_uint32),
        ("param_p", C.c_float),
        ("param_q", C.c_float),
        ("init_strategy", C.c_uint32),
        ("search_mode", C.c_uint32),
        ("seed", C.c_uint64),
        ("device", C.c_int32),
        ("rank", C.c_uint32),
        ("world", C.c_uint32),
        ("generate_mode", C.c_uint32),
    ]


class WmhMemory(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("graph_bytes", "graph_pool_bytes", "walks_bytes", "index_bytes", "scratch_bytes", "total_bytes")]


class WharfError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"wharfmh_b200 error {code}: {message}")
        self.code = code


# name -> (restype, argtypes); every symbol include/wharfmh_b200.h declares
u64, u32, i32, vp, cp = C.c_uint64, C.c_uint32, C.c_int, C.c_void_p, C.c_char_p
PU64, PU32 = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
SIGNATURES = {
    "wmh_default_config": (None, [C.POINTER(WmhConfig)]),
    "wmh_last_error": (cp, []),
    "wmh_version": (cp, []),
    "wmh_create": (i32, [C.POINTER(vp), u64, u64, vp, vp, C.POINTER(WmhConfig)]),
    "wmh_create_device": (i32, [C.POINTER(vp), u64, u64, vp, vp, C.POINTER(WmhConfig)]),
    "wmh_create_empty": (i32, [C.POINTER(vp), u64, C.POINTER(WmhConfig)]),
    "wmh_destroy": (None, [vp]),
    "wmh_destroy_index": (i32, [vp]),
    "wmh_set_config": (i32, [vp, C.POINTER(WmhConfig)]),
    "wmh_get_config": (i32, [vp, C.POINTER(WmhConfig)]),
    "wmh_number_of_vertices": (u64, [vp]),
    "wmh_number_of_edges": (u64, [vp]),
    "wmh_flatten_graph": (i32, [vp, vp, vp]),
    "wmh_degree": (i32, [vp, u32, PU64]),
    "wmh_neighbors": (i32, [vp, u32, vp, u64, PU64]),
    "wmh_generate_initial_random_walks": (i32, [vp]),
    "wmh_generate_initial_random_walks_to_host": (i32, [vp, vp, u64]),
    "wmh_last_steps": (u64, [vp]),
    "wmh_number_of_walks": (u64, [vp]),
    "wmh_number_of_local_walks": (u64, [vp]),
    "wmh_insert_edges_batch": (i32, [vp, u64, vp, i32, i32, u64, i32, PU64]),
    "wmh_delete_edges_batch": (i32, [vp, u64, vp, i32, i32, u64, i32, PU64]),
    "wmh_insert_edges_batch_device": (i32, [vp, u64, vp, i32, i32, PU64]),
    "wmh_delete_edges_batch_device": (i32, [vp, u64, vp, i32, i32, PU64]),
    "wmh_affected_walks": (i32, [vp, vp, vp, u64]),
    "wmh_affected_walks_device": (i32, [vp, C.POINTER(vp), C.POINTER(vp), PU64]),
    "wmh_affected_walks_global_device": (i32, [vp, vp, u64]),
    "wmh_sampler_chain": (i32, [vp, u32, u32, u64, PU32, vp, u64]),
    "wmh_touched_vertices": (i32, [vp, vp, u64, PU64]),
    "wmh_walk": (i32, [vp, u64, vp, PU32]),
    "wmh_walk_string": (i32, [vp, u64, vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "wmh_vertex_at_walk": (i32, [vp, u64, u32, PU32]),
    "wmh_walks_copy": (i32, [vp, u64, u64, vp]),
    "wmh_walks_gather": (i32, [vp, vp, u64, vp]),
    "wmh_local_to_global_walk": (u64, [vp, u64]),
    "wmh_global_to_local_walk": (C.c_int64, [vp, u64]),
    "wmh_walk_owner": (u32, [vp, u64]),
    "wmh_local_walk_ids": (u64, [vp, vp, u64]),
    "wmh_random_stream": (None, [u64, u64, vp]),
    "wmh_unit_double": (C.c_double, [u64]),
    "wmh_keyed_draw": (u64, [u64, u64, u64, u64, u64]),
    "wmh_partition_owner": (u32, [u64, u32, u32, u64]),
    "wmh_partition_local_walk_ids": (u64, [u64, u32, u32, u32, vp, u64]),
    "wmh_parse_adjacency_graph": (i32, [vp, u64, PU64, PU64, vp, vp]),
    "wmh_format_adjacency_graph": (u64, [u64, u64, vp, vp, vp, u64]),
    "wmh_parse_edge_list": (i32, [vp, u64, vp, u64, PU64, PU64]),
    "wmh_format_walks": (u64, [vp, u64, u32, vp, u64]),
    "wmh_write_walks": (i32, [vp, vp, u64, cp, i32, PU64]),
    "wmh_corpus_fingerprint": (i32, [vp, PU64]),
    "wmh_build_index": (i32, [vp]),
    "wmh_index_find_next": (i32, [vp, u32, u64, u32, PU32, C.POINTER(i32)]),
    "wmh_index_size": (i32, [vp, u32, PU64]),
    "wmh_index_sizes": (i32, [vp, vp]),
    "wmh_szudzik_pair_u64": (i32, [i32, vp, vp, vp, u64, PU64]),
    "wmh_szudzik_unpair_u64": (i32, [i32, vp, vp, vp, u64]),
    "wmh_szudzik_pair_u32": (i32, [i32, vp, vp, vp, u64, PU64]),
    "wmh_szudzik_unpair_u32": (i32, [i32, vp, vp, vp, u64]),
    "wmh_szudzik_pair_triplet_u64": (i32, [i32, vp, vp, vp, vp, u64, PU64]),
    "wmh_szudzik_unpair_triplet_u64": (i32, [i32, vp, vp, vp, vp, u64]),
    "wmh_szudzik_pair": (u64, [u64, u64]),
    "wmh_szudzik_unpair": (None, [u64, PU64, PU64]),
    "wmh_compressed_walks": (i32, [vp, vp, vp, u64, PU64]),
    "wmh_compressed_find_next": (i32, [vp, u32, u64, u32, PU32, C.POINTER(i32)]),
    "wmh_timer_get": (C.c_double, [vp, i32]),
    "wmh_timer_reset": (None, [vp, i32]),
    "wmh_memory_footprint": (i32, [vp, C.POINTER(WmhMemory)]),
    "wmh_kernel_launches": (u64, [vp]),
    "wmh_last_walk_kernel_ms": (C.c_double, [vp]),
    "wmh_last_index_ms": (C.c_double, [vp]),
    "wmh_last_search_mode": (u32, [vp]),
    "wmh_neighbour_filter": (u32, [vp]),
    "wmh_filter_check": (i32, [vp, PU64]),
    "wmh_scan_touched_no_alloc": (i32, [vp, vp, u64, vp]),
    "wmh_device_walks": (vp, [vp]),
    "wmh_stream": (vp, [vp]),
    "wmh_synchronize": (i32, [vp]),
    "wmh_trim_memory": (i32, [i32]),
    "wmh_peer_create": (i32, [vp, u64, vp]),
    "wmh_peer_attach": (i32, [vp, vp]),
    "wmh_peer_detach": (i32, [vp]),
    "wmh_peer_bytes_needed": (u64, [vp]),
    "wmh_peer_exchanges": (u64, [vp]),
    "wmh_generate_batch_of_edges_device": (i32, [i32, u64, u64, u64, i32, i32, C.c_double, C.c_double, C.c_double, vp, PU64]),
    "wmh_generate_batch_of_edges": (i32, [i32, u64, u64, u64, i32, i32, C.c_double, C.c_double, C.c_double, vp, u64, PU64]),
    "wmh_pairs_to_csr_device": (i32, [i32, u64, vp, u64, vp, vp]),
}

_lib = None


def load() -> C.CDLL:
    """Loads libwharfmh_b200.so; raises when it is absent (build with __graft_entry__.build())."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("WHARFMH_B200_LIB", LIB_PATH))
        if not path.exists():
            raise ImportError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              f"or `make -C {PKG_DIR / 'csrc'}`; there is no fallback implementation")
        lib = C.CDLL(str(path))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)          # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != OK:
        raise WharfError(rc, load().wmh_last_error().decode("utf-8", "replace"))
