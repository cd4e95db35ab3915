"""B200-native WharfMH hot path: bulk random-walk generation and edge-batch walk maintenance.

Public surface = the reference's (dygrl::WharfMH + namespace config), see wharfmh.py.  All compute is CUDA
(sm_100a) behind the C ABI of include/wharfmh_b200.h; importing this package without the built library
raises.
"""
from . import _capi, graph_io
from ._capi import (DEEPWALK, GENERATE_AUTO, GENERATE_COALESCED, GENERATE_PER_WALKER, INIT_BURNIN, INIT_RANDOM, INIT_WEIGHT, NODE2VEC, NO_VERTEX, REF_SENTINEL, SEARCH_AUTO,
                    SEARCH_INDEX, SEARCH_SCAN, WharfError)
from .wharfmh import (Config, WharfMH, generate_batch_of_edges, keyed_draw, partition_local_walk_ids, partition_owner, random_stream, szudzik_pair, szudzik_unpair, trim_memory,
                      unit_double)

_capi.load()   # fail loudly at import when the CUDA library is missing

__all__ = ["Config", "WharfMH", "WharfError", "DEEPWALK", "NODE2VEC", "INIT_RANDOM", "INIT_BURNIN", "INIT_WEIGHT",
           "SEARCH_AUTO", "SEARCH_INDEX", "SEARCH_SCAN", "GENERATE_AUTO", "GENERATE_PER_WALKER", "GENERATE_COALESCED", "NO_VERTEX", "REF_SENTINEL", "szudzik_pair", "szudzik_unpair", "trim_memory", "partition_owner", "partition_local_walk_ids", "graph_io", "random_stream", "unit_double", "keyed_draw", "generate_batch_of_edges"]
