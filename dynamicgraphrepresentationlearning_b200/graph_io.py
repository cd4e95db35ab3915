"""Ingest and export: the callers either side of the path.

What the reference's drivers do before they construct a WharfMH and after a batch, with the same names where the
reference has one:

  read_unweighted_graph      libs/compressed_trees/common/IO.h:67-106 (Ligra "AdjacencyGraph" text)
  write_unweighted_graph     the file experiments/bin/SNAPtoAdj writes
  read_edge_list             SNAP "u v" lines: SNAPtoAdj's input, create_edge_stream's input
  snap_to_adj                SNAPtoAdj -s: symmetrise, drop loops and duplicates, ascending rows -> CSR
  graph_from_edge_list       the same, left on the device as an engine
  create_edge_stream         experiments/src/vertex-classification.cpp:5-38
  write_walks                the walks.txt hand-off of vertex-classification.cpp:167-178

Text parsing and formatting are host functions of the C library (wmh_parse_*, wmh_format_*); the set work of ingest
(symmetrise + sort + de-duplicate + per-row merge) is the engine's own batch insert on the device - an empty
WharfMH(n) receives the edge list as one batch with remove_dups, exactly how the reference's tests grow a graph.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _capi
from ._capi import check
from .wharfmh import Config, WharfMH, _ptr


def _text_of(source) -> bytes:
    if isinstance(source, (bytes, bytearray, memoryview)):
        return bytes(source)
    with open(os.fspath(source), "rb") as f:
        return f.read()


def read_unweighted_graph(source, is_symmetric: bool = True, mmap: bool = False):
    """(n, m, offsets[n] uint64, edges[m] uint32) of a Ligra AdjacencyGraph file (or its bytes).  As in the reference
    there is no n+1-th offset; is_symmetric and mmap are accepted for signature parity and unused (IO.h:67)."""
    lib = _capi.load()
    text = _text_of(source)
    n, m = C.c_uint64(), C.c_uint64()
    check(lib.wmh_parse_adjacency_graph(text, len(text), C.byref(n), C.byref(m), None, None))
    offsets = np.empty(n.value, dtype=np.uint64)
    edges = np.empty(m.value, dtype=np.uint32)
    check(lib.wmh_parse_adjacency_graph(text, len(text), C.byref(n), C.byref(m), _ptr(offsets), _ptr(edges)))
    return int(n.value), int(m.value), offsets, edges


def format_unweighted_graph(offsets, edges) -> bytes:
    lib = _capi.load()
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    edges = np.ascontiguousarray(edges, dtype=np.uint32)
    n, m = offsets.shape[0], edges.shape[0]
    need = lib.wmh_format_adjacency_graph(n, m, _ptr(offsets), _ptr(edges), None, 0)
    buf = C.create_string_buffer(need)
    lib.wmh_format_adjacency_graph(n, m, _ptr(offsets), _ptr(edges), buf, need)
    return buf.raw


def write_unweighted_graph(path, offsets, edges, row_ptr: bool = False) -> None:
    """Writes the AdjacencyGraph file.  offsets holds n entries as in the reference; pass row_ptr=True for the
    n+1-entry row_ptr flatten_graph() returns (its last entry is dropped)."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    with open(os.fspath(path), "wb") as f:
        f.write(format_unweighted_graph(offsets[:-1] if row_ptr else offsets, edges))


def read_edge_list(source) -> tuple[np.ndarray, int]:
    """(pairs [k, 2] uint32 in file order, vertices = largest id + 1) of a SNAP edge list (path or bytes)."""
    lib = _capi.load()
    text = _text_of(source)
    k, vertices = C.c_uint64(), C.c_uint64()
    check(lib.wmh_parse_edge_list(text, len(text), None, 0, C.byref(k), C.byref(vertices)))
    pairs = np.empty((k.value, 2), dtype=np.uint32)
    check(lib.wmh_parse_edge_list(text, len(text), _ptr(pairs), k.value, C.byref(k), C.byref(vertices)))
    return pairs, int(vertices.value)


def symmetrized(pairs) -> np.ndarray:
    """(u, v), (v, u) for every input pair, interleaved - the order create_edge_stream pushes them in."""
    pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    return np.stack([pairs, pairs[:, ::-1]], axis=1).reshape(-1, 2)


def graph_from_edge_list(pairs, vertices: int | None = None, symmetric: bool = True, config: Config | None = None) -> WharfMH:
    """An engine holding the graph of an edge list: loops and duplicate edges dropped, rows ascending; with
    symmetric=True every edge in both directions (SNAPtoAdj -s).  Built on the device: an empty WharfMH(n)
    (graph/wharfmh.h:26) takes the list as one insert batch with remove_dups (graph/wharfmh.h:439-470)."""
    pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    if vertices is None:
        vertices = int(pairs.max()) + 1 if pairs.size else 0
    batch = symmetrized(pairs) if symmetric else pairs.copy()
    engine = WharfMH(vertices, int(batch.shape[0]), config=config)
    try:
        if batch.shape[0]:
            engine.insert_edges_batch(batch, sorted=False, remove_dups=True, apply_walk_updates=False)
    except Exception:
        engine.destroy()
        raise
    return engine


def snap_to_adj(source, symmetric: bool = True, device: int = 0):
    """SNAPtoAdj [-s]: edge list (path, bytes or a [k, 2] array) -> (n, m, offsets[n], edges[m])."""
    if isinstance(source, np.ndarray):
        pairs, vertices = source, None
    else:
        pairs, vertices = read_edge_list(source)
    if np.asarray(pairs).size == 0:
        return 0, 0, np.empty(0, dtype=np.uint64), np.empty(0, dtype=np.uint32)
    with graph_from_edge_list(pairs, vertices, symmetric, Config(device=device)) as engine:
        row_ptr, col = engine.flatten_graph()
        return engine.number_of_vertices(), int(col.shape[0]), row_ptr[:-1].copy(), col


def create_edge_stream(source, edge_partition_size: int = 5000) -> list[np.ndarray]:
    """vertex-classification.cpp:5-38: the file's lines in chunks of edge_partition_size, every line contributing
    (u, v) and (v, u); one [2k, 2] uint32 batch per chunk, ready for insert_edges_batch(..., remove_dups=True)."""
    if edge_partition_size <= 0:
        raise ValueError("edge_partition_size must be positive")
    pairs = source if isinstance(source, np.ndarray) else read_edge_list(source)[0]
    pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    return [symmetrized(pairs[i: i + edge_partition_size]) for i in range(0, pairs.shape[0], edge_partition_size)]


def format_walks(rows) -> bytes:
    """Rows of vertices (NO_VERTEX padded) as the text of WharfMH::walk, one walk per line."""
    lib = _capi.load()
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    if rows.ndim != 2:
        raise ValueError("rows must be [walks, walk_length]")
    need = lib.wmh_format_walks(_ptr(rows), rows.shape[0], rows.shape[1], None, 0)
    buf = np.empty(need, dtype=np.uint8)
    lib.wmh_format_walks(_ptr(rows), rows.shape[0], rows.shape[1], _ptr(buf), need)
    return buf.tobytes()


def write_walks(engine: WharfMH, path, walk_ids, append: bool = False) -> int:
    """walks.txt (vertex-classification.cpp:167-178): the named walks (global ids this engine holds, e.g. what
    insert_edges_batch returned) gathered from the device and written as text.  Returns the bytes written."""
    return engine.write_walks(path, walk_ids, append=append)


def main(argv=None) -> int:
    """`python -m dynamicgraphrepresentationlearning_b200.graph_io [-s] [-f] <edge list> <adjacency graph>`: the command
    line of experiments/bin/SNAPtoAdj (-s symmetrises; -f is accepted and ignored as there), with the set work done on
    the device."""
    import sys
    args = [a for a in (sys.argv[1:] if argv is None else argv)]
    symmetric = "-s" in args
    files = [a for a in args if a not in ("-s", "-f")]
    if len(files) != 2:
        print("usage: graph_io [-s] [-f] <SNAP edge list> <AdjacencyGraph output>", file=sys.stderr)
        return 2
    n, m, offsets, edges = snap_to_adj(files[0], symmetric=symmetric)
    write_unweighted_graph(files[1], offsets, edges)
    print(f"Vertices: {n} Edges: {m}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
