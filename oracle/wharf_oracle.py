"""TEST INFRASTRUCTURE ONLY: ctypes front-end of oracle/wharf_oracle.c and a runner for oracle/_ref/wharf_ref.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product package (dynamicgraphrepresentationlearning_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "_build" / "libwharf_oracle.so"
REF_BIN = HERE / "_ref" / "wharf_ref"
NO_VERTEX = 0xFFFFFFFF

DEEPWALK, NODE2VEC = 0, 1
INIT_RANDOM, INIT_BURNIN, INIT_WEIGHT = 0, 1, 2


class WoConfig(C.Structure):
    _fields_ = [
        ("walks_per_vertex", C.c_uint32),
        ("walk_length", C.c_uint32),
        ("model", C.c_uint32),
        ("deterministic", C.c_uint32),
        ("param_p", C.c_float),
        ("param_q", C.c_float),
        ("init_strategy", C.c_uint32),
        ("reserved", C.c_uint32),
        ("seed", C.c_uint64),
    ]


def build(force: bool = False) -> Path:
    """Compiles the C restatement (and the real reference when /root/reference is present)."""
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < (HERE / "wharf_oracle.c").stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "oracle"], check=True, capture_output=True)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(LIB_PATH))
        u64, u32, p = C.c_uint64, C.c_uint32, C.c_void_p
        L.wo_random_seed.argtypes = [u64, C.POINTER(u64)]
        L.wo_lrand.argtypes = [C.POINTER(u64)]
        L.wo_lrand.restype = u64
        L.wo_drand.argtypes = [C.POINTER(u64)]
        L.wo_drand.restype = C.c_double
        L.wo_draw_table.argtypes = [u32, u32, p]
        for name, t in (("32", u32), ("64", u64)):
            getattr(L, "wo_pair" + name).argtypes = [t, t]
            getattr(L, "wo_pair" + name).restype = t
            getattr(L, "wo_unpair" + name).argtypes = [t, C.POINTER(t), C.POINTER(t)]
            getattr(L, "wo_pair_triplet" + name).argtypes = [t, t, t]
            getattr(L, "wo_pair_triplet" + name).restype = t
            getattr(L, "wo_unpair_triplet" + name).argtypes = [t, C.POINTER(t), C.POINTER(t), C.POINTER(t)]
        L.wo_compressed_find_next.argtypes = [p, u64, u64, u32, u32]
        L.wo_compressed_find_next.restype = u32
        L.wo_hash32.argtypes = [u32]
        L.wo_hash32.restype = u32
        L.wo_hash64.argtypes = [u64]
        L.wo_hash64.restype = u64
        L.wo_keyed.argtypes = [u64] * 5
        L.wo_keyed.restype = u64
        L.wo_generate_batch.argtypes = [u64, u64, u64, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, p]
        L.wo_generate_batch.restype = u64
        L.wo_create.argtypes = [u64, u64, p, p, C.POINTER(WoConfig)]
        L.wo_create.restype = p
        L.wo_destroy.argtypes = [p]
        L.wo_generate.argtypes = [p]
        L.wo_generate.restype = u64
        L.wo_update.argtypes = [p, p, u64, C.c_int, C.c_int, C.c_int]
        L.wo_update.restype = C.c_int
        for f in ("wo_num_vertices", "wo_num_edges", "wo_num_affected", "wo_steps_last"):
            getattr(L, f).argtypes = [p]
            getattr(L, f).restype = u64
        for f in ("wo_row_ptr", "wo_col", "wo_walks", "wo_affected", "wo_affected_pos"):
            getattr(L, f).argtypes = [p]
            getattr(L, f).restype = p
        L.wo_sampler_init_value.argtypes = [p, u32, u32]
        L.wo_sampler_init_value.restype = u32
        L.wo_sampler_chain.argtypes = [p, u32, u32, u64, p]
        L.wo_sampler_chain.restype = u32
        L.wo_corpus_fnv1a.argtypes = [p, u64, u32]
        L.wo_corpus_fnv1a.restype = u64
        L.wo_corpus_fingerprint.argtypes = [p, u64, u32]
        L.wo_corpus_fingerprint.restype = u64
        L.wo_threads.restype = C.c_int
        _lib = L
    return _lib


def _ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def random_stream(seed: int, k: int) -> list[int]:
    s = (C.c_uint64 * 2)()
    lib().wo_random_seed(seed, s)
    return [int(lib().wo_lrand(s)) for _ in range(k)]


def drand_stream(seed: int, k: int) -> list[float]:
    s = (C.c_uint64 * 2)()
    lib().wo_random_seed(seed, s)
    return [float(lib().wo_drand(s)) for _ in range(k)]


def draw_table(replicas: int, length: int) -> np.ndarray:
    out = np.empty((replicas, length), dtype=np.uint64)
    lib().wo_draw_table(replicas, length, _ptr(out))
    return out


def generate_batch(edges_number: int, vertices_number: int, seed: int, self_loops: bool = False,
                   directed: bool = True, a: float = 0.5, b: float = 0.2, c: float = 0.1) -> np.ndarray:
    """utility::generate_batch_of_edges -> (k, 2) uint32, sorted by (src, dst), deduplicated."""
    total = edges_number if directed else 2 * edges_number
    out = np.empty((max(total, 1), 2), dtype=np.uint32)
    k = lib().wo_generate_batch(edges_number, vertices_number, seed, int(self_loops), int(directed), a, b, c, _ptr(out))
    return out[:k].copy()


def rmat_graph(scale: int, edges_half: int, seed: int = 42) -> tuple[np.ndarray, np.ndarray]:
    """Symmetric RMAT CSR as SURVEY.md 8d prescribes: generate_batch_of_edges(E/2, 2^s + 1, seed,
    self_loops=false, directed=false) -> (offsets[n] u64 (no n+1-th), edges u32), n = 2^scale."""
    n = 1 << scale
    pairs = generate_batch(edges_half, n + 1, seed, False, False)
    return csr_from_sorted_pairs(n, pairs)


def csr_from_sorted_pairs(n: int, pairs: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    deg = np.bincount(pairs[:, 0].astype(np.int64), minlength=n).astype(np.uint64)
    offsets = np.zeros(n, dtype=np.uint64)
    if n > 1:
        np.cumsum(deg[:-1], out=offsets[1:])
    return offsets, np.ascontiguousarray(pairs[:, 1]).astype(np.uint32)


def corpus_hash(walks: np.ndarray) -> int:
    walks = np.ascontiguousarray(walks, dtype=np.uint32)
    return int(lib().wo_corpus_fnv1a(_ptr(walks), walks.shape[0], walks.shape[1]))


def corpus_fingerprint(walks: np.ndarray) -> int:
    """Wrapping sum of the per-walk FNV-1a of the text form: what wmh_corpus_fingerprint computes on the device."""
    walks = np.ascontiguousarray(walks, dtype=np.uint32)
    return int(lib().wo_corpus_fingerprint(_ptr(walks), walks.shape[0], walks.shape[1]))


class Oracle:
    """Plain-C restatement of dygrl::WharfMH (graph/wharfmh.h:21) for one graph + config."""

    def __init__(self, n: int, offsets: np.ndarray, edges: np.ndarray, *, walks_per_vertex: int = 10,
                 walk_length: int = 80, model: int = DEEPWALK, p: float = 0.5, q: float = 2.0,
                 deterministic: bool = True, init_strategy: int = INIT_WEIGHT, seed: int = 0):
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)[:n]
        edges = np.ascontiguousarray(edges, dtype=np.uint32)
        self.cfg = WoConfig(walks_per_vertex, walk_length, model, int(deterministic), p, q, init_strategy, 0, seed)
        self.L = walk_length
        self.wpv = walks_per_vertex
        self._h = lib().wo_create(n, edges.shape[0], _ptr(offsets), _ptr(edges), C.byref(self.cfg))

    def close(self):
        if self._h:
            lib().wo_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def n(self) -> int:
        return int(lib().wo_num_vertices(self._h))

    @property
    def m(self) -> int:
        return int(lib().wo_num_edges(self._h))

    def generate(self) -> int:
        return int(lib().wo_generate(self._h))

    def update(self, pairs: np.ndarray, delete: bool = False, remove_dups: bool = True, apply_walk_updates: bool = True) -> int:
        pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        rc = lib().wo_update(self._h, _ptr(pairs), pairs.shape[0], int(delete), int(remove_dups), int(apply_walk_updates))
        if rc != 0:
            raise ValueError("vertex id out of range")
        return int(lib().wo_num_affected(self._h))

    def _view(self, fn, count, dtype):
        addr = getattr(lib(), fn)(self._h)
        if not addr or count == 0:
            return np.empty(0, dtype=dtype)
        buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(addr)
        return np.frombuffer(buf, dtype=dtype, count=count).copy()

    def walks(self) -> np.ndarray:
        W = self.n * self.wpv
        return self._view("wo_walks", W * self.L, np.uint32).reshape(W, self.L)

    def csr(self) -> tuple[np.ndarray, np.ndarray]:
        return self._view("wo_row_ptr", self.n + 1, np.uint64), self._view("wo_col", self.m, np.uint32)

    def affected(self) -> tuple[np.ndarray, np.ndarray]:
        k = int(lib().wo_num_affected(self._h))
        return self._view("wo_affected", k, np.uint64), self._view("wo_affected_pos", k, np.uint32)

    def steps_last(self) -> int:
        return int(lib().wo_steps_last(self._h))

    def sampler_init_value(self, cur: int, aux: int) -> int:
        return int(lib().wo_sampler_init_value(self._h, cur, aux))

    def sampler_chain(self, cur: int, aux: int, iters: int) -> tuple[int, np.ndarray]:
        """MetropolisHastingsSampler::sample as a chain on the fixed state (cur, aux): (init value, counts per neighbour slot)."""
        row_ptr = self.csr()[0]
        degree = int(row_ptr[cur + 1] - row_ptr[cur])
        counts = np.zeros(max(degree, 1), dtype=np.uint64)
        init = int(lib().wo_sampler_chain(self._h, cur, aux, iters, _ptr(counts)))
        return init, counts[:degree]


# ---------------------------------------------------------------------------------------------------
# File formats of oracle/ref_driver.cpp and a runner for the compiled, unmodified reference.
# ---------------------------------------------------------------------------------------------------

def write_graph(path, n: int, offsets: np.ndarray, edges: np.ndarray) -> None:
    with open(path, "wb") as f:
        np.array([n, edges.shape[0]], dtype=np.uint64).tofile(f)
        np.ascontiguousarray(offsets, dtype=np.uint64)[:n].tofile(f)
        np.ascontiguousarray(edges, dtype=np.uint32).tofile(f)


def write_batch(path, pairs: np.ndarray) -> None:
    pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    with open(path, "wb") as f:
        np.array([pairs.shape[0]], dtype=np.uint64).tofile(f)
        pairs.tofile(f)


def read_batch(path) -> np.ndarray:
    with open(path, "rb") as f:
        k = int(np.fromfile(f, dtype=np.uint64, count=1)[0])
        return np.fromfile(f, dtype=np.uint32, count=2 * k).reshape(k, 2)


def read_walks(path) -> np.ndarray:
    with open(path, "rb") as f:
        W, L = (int(x) for x in np.fromfile(f, dtype=np.uint64, count=2))
        return np.fromfile(f, dtype=np.uint32, count=W * L).reshape(W, L)


def read_ids(path) -> np.ndarray:
    with open(path, "rb") as f:
        k = int(np.fromfile(f, dtype=np.uint64, count=1)[0])
        return np.fromfile(f, dtype=np.uint64, count=k)


def ref_available() -> bool:
    return REF_BIN.exists() and os.access(REF_BIN, os.X_OK)


def ref_run(args: list[str], threads: int | None = None, timeout: float | None = None) -> list[dict]:
    """Runs oracle/_ref/wharf_ref (the unmodified reference) and parses its JSON lines."""
    env = dict(os.environ)
    if threads is not None:
        env["NUM_THREADS"] = str(threads)       # pbbslib/scheduler.h:210
    out = subprocess.run([str(REF_BIN)] + [str(a) for a in args], check=True, capture_output=True, text=True,
                         env=env, timeout=timeout)
    lines = []
    for ln in out.stdout.splitlines():
        ln = ln.strip()
        if ln.startswith("{"):
            lines.append(json.loads(ln))
    return lines
