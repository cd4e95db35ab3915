"""CPU tests of the drop-in boundary: the shared library loads and exports every symbol include/wharfmh_b200.h
declares; host-only entry points behave; compute entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
HEADER = ROOT / "include" / "wharfmh_b200.h"


def declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"WMH_API\s+[\w\s\*]+?\b(wmh_\w+)\s*\(", text)))


def test_header_declares_the_reference_surface():
    syms = declared_symbols()
    for must in ("wmh_create", "wmh_generate_initial_random_walks", "wmh_insert_edges_batch", "wmh_delete_edges_batch",
                 "wmh_walk", "wmh_vertex_at_walk", "wmh_flatten_graph", "wmh_destroy", "wmh_destroy_index",
                 "wmh_memory_footprint", "wmh_szudzik_pair_u64", "wmh_index_find_next", "wmh_compressed_find_next"):
        assert must in syms
    # every entry point cites the reference interface it replaces
    text = HEADER.read_text()
    for ref in ("graph/wharfmh.h:58", "graph/wharfmh.h:250", "graph/wharfmh.h:439", "graph/wharfmh.h:365", "walks/pairings.h",
                "walks/inverted_index.h:74-92", "walks/compressed_walks.h", "config/globals.h:7-29"):
        assert ref in text


def test_library_exports_every_declared_symbol(wharf):
    lib = wharf._capi.load()
    syms = declared_symbols()
    assert len(syms) >= 50
    for name in syms:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in wharf._capi.SIGNATURES, f"{name} has no ctypes signature"
    assert set(wharf._capi.SIGNATURES) == set(syms)


def test_host_only_entry_points(wharf, golden):
    lib = wharf._capi.load()
    assert b"sm_100a" in lib.wmh_version()
    cfg = wharf._capi.WmhConfig()
    lib.wmh_default_config(C.byref(cfg))
    # config/globals.h:7-29 defaults
    assert (cfg.walks_per_vertex, cfg.walk_length, cfg.model, cfg.deterministic) == (10, 80, wharf.NODE2VEC, 1)
    assert (cfg.param_p, cfg.param_q, cfg.init_strategy, cfg.world) == (4.0, 1.0, wharf.INIT_WEIGHT, 1)
    # Szudzik host scalars: tests/pairings.cpp KATs and the reference's own outputs
    assert wharf.szudzik_pair(65535, 65535) == 4294967295 and wharf.szudzik_unpair(4294967295) == (65535, 65535)
    for x, y, z in golden["pair64"]:
        assert wharf.szudzik_pair(x, y) == z and wharf.szudzik_unpair(z) == (x, y)
    for z, x, y in golden["unpair64"]:
        assert wharf.szudzik_unpair(z) == (x, y)
    # exact above the reference's double-sqrt domain: z = s*s for large s decodes to (s, 0)
    s = 2**32 - 3
    assert wharf.szudzik_unpair(s * s) == (s, 0) and wharf.szudzik_unpair(s * s + s) == (0, s)
    rng = np.random.default_rng(0)
    for x, y in rng.integers(0, 2**32, (5000, 2), dtype=np.uint64):
        assert wharf.szudzik_unpair(wharf.szudzik_pair(int(x), int(y))) == (int(x), int(y))


@pytest.mark.parametrize("n,world,wpv", [(512, 2, 2), (509, 3, 2), (1000, 8, 10), (7, 8, 3), (64, 5, 10), (300, 4, 10), (129, 2, 5),
                                         (1 << 14, 8, 10), (1 << 14, 7, 10), (1000, 6, 9), (33, 1, 4), (10, 8, 1)])
def test_partition_rule_on_the_host(wharf, n, world, wpv):
    """The sharding rule (whole replicas first, the remaining replicas in gcd(world, rest) groups by rotated start-vertex
    slots) is host arithmetic: every walk has exactly one owner, local order is ascending, shares are balanced, and
    wmh_partition_owner agrees with the id lists.  (On a GPU tests/test_gpu_parity.py::test_shard_rule_is_a_partition
    checks the engine against the same lists.)"""
    import math
    owner = np.full(n * wpv, -1, dtype=np.int64)
    q = wpv // world if world > 1 else 0
    rest = wpv - world * q if world > 1 else wpv
    g = math.gcd(world, rest) if rest else world
    counts = []
    for r in range(world):
        ids = wharf.partition_local_walk_ids(n, wpv, r, world).astype(np.int64)
        assert ids.size == 0 or ((np.diff(ids) > 0).all() and ids.min() >= 0 and ids.max() < n * wpv)
        assert (owner[ids] == -1).all()
        owner[ids] = r
        counts.append(ids.size)
        if q:
            assert np.array_equal(ids[: q * n], np.arange(r * q * n, (r + 1) * q * n))      # its whole replicas
        for walk in ids[:: max(1, ids.size // 97)]:
            assert wharf.partition_owner(n, wpv, world, int(walk)) == r
    assert (owner >= 0).all()
    shared_per_rank = (rest // g) * n // (world // g)
    assert max(counts) - min(counts) <= max(1, rest // g) and abs(counts[0] - (q * n + shared_per_rank)) <= max(1, rest // g)
    # a shared replica's start vertices rotate from block to block inside its group
    if world == 8 and n >= 1000 and wpv == 10:
        shared = owner[8 * n: 9 * n]
        assert set(shared.tolist()) == {0, 1, 2, 3} and len({int(shared[b * 4]) for b in range(100)}) > 1


def test_generators_on_the_host(wharf, wo, golden):
    """SURVEY 8a row a3 pinned on the PRODUCT's own host functions (the code the draw table and the kernels are built
    from, csrc/rng.cuh), without a GPU: utility::Random's stream against the unmodified reference's values
    (tests/golden/golden.json "random", "drand"), the keyed generator against the CPU restatement's."""
    assert wharf.random_stream(0, 3).tolist() == [10407335079877134008, 3962074050977524353, 12330921719341810270]
    assert wharf.random_stream(1, 3).tolist() == [7437716154062571297, 1116485288419732103, 9741853272534875990]
    assert wharf.random_stream(2, 3).tolist() == [2528678525309737254, 8345201964650685463, 16366336586872171348]
    for seed, values in golden["random"].items():
        assert wharf.random_stream(int(seed), len(values)).tolist() == values
    for seed, values in golden["drand"].items():                        # drand = unit_double(lrand()), utils/utility.h:208-218
        assert [wharf.unit_double(int(x)) for x in wharf.random_stream(int(seed), len(values))] == values
    table = wo.draw_table(10, 80)                                       # deterministic mode's only randomness: X[r][k]
    for r in range(10):
        assert np.array_equal(wharf.random_stream(r, 80), table[r])
    assert wharf.unit_double(0) == 0.0 and wharf.unit_double(2 ** 64 - 1) == 1.0 - 2.0 ** -52
    rng = np.random.default_rng(11)
    L = wo.lib()
    for _ in range(2000):
        seed, a, b, c = (int(x) for x in rng.integers(0, 2 ** 63, size=4, dtype=np.uint64) * 2 + rng.integers(0, 2, size=4, dtype=np.uint64))
        dom = int(rng.integers(1, 4))
        assert wharf.keyed_draw(seed, dom, a, b, c) == int(L.wo_keyed(seed, dom, a, b, c))
    assert wharf.keyed_draw(1, 1, 0, 0, 0) != wharf.keyed_draw(1, 1, 0, 0, 1) != wharf.keyed_draw(1, 2, 0, 0, 0)


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: include/wharfmh_b200.h compiles as C99 (pedantic, warnings as errors) and a C program
    links the library and calls host-only entry points."""
    import subprocess
    lib_dir = ROOT / "dynamicgraphrepresentationlearning_b200" / "lib"
    src = tmp_path / "c_abi.c"
    src.write_text('#include "wharfmh_b200.h"\n#include <stdio.h>\n'
                   'int main(void) { wmh_config c; uint64_t x = 0, y = 0, s[2];\n'
                   '  wmh_default_config(&c); wmh_szudzik_unpair(wmh_szudzik_pair(123, 25), &x, &y); wmh_random_stream(0, 2, s);\n'
                   '  printf("%u %u %llu %llu %llu %u\\n", c.walks_per_vertex, c.walk_length, (unsigned long long) x, (unsigned long long) y,\n'
                   '         (unsigned long long) s[0], wmh_partition_owner(100, 10, 2, 999));\n'
                   '  return wmh_last_error() == 0; }\n')
    exe = tmp_path / "c_abi"
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", f"-I{ROOT / 'include'}", str(src), "-o", str(exe),
                    f"-L{lib_dir}", "-lwharfmh_b200", f"-Wl,-rpath,{lib_dir}"], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    assert out == ["10", "80", "123", "25", "10407335079877134008", "1"]       # config/globals.h defaults; pairing KAT; Random(0); replica 9 of 10 -> rank 1


def test_no_cpu_fallback(wharf):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(wharf.WharfError) as ei:
        wharf.WharfMH(4, 0, np.zeros(4, dtype=np.uint64), np.zeros(0, dtype=np.uint32))
    assert ei.value.code == wharf._capi.ERR_CUDA and "no CPU path" in str(ei.value)
    lib = wharf._capi.load()
    k = C.c_uint64()
    assert lib.wmh_generate_batch_of_edges_device(0, 10, 17, 0, 0, 0, 0.5, 0.2, 0.1, C.c_void_p(8), C.byref(k)) == wharf._capi.ERR_CUDA
    assert lib.wmh_generate_initial_random_walks(None) == wharf._capi.ERR_INVALID


def test_product_never_touches_the_oracle():
    pkg = ROOT / "dynamicgraphrepresentationlearning_b200"
    for path in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.hpp")) + [ROOT / "include" / "wharfmh_b200.h"]:
        text = path.read_text()
        assert "wharf_oracle" not in text and "oracle/" not in text and "from oracle" not in text, path
