"""Compiles the reference's own callers of the boundary - unchanged, from where they lie under /root/reference - against
include/compat + include/wharfmh_b200.hpp and links them with libwharfmh_b200.so.  Outputs only into
tests/cpp/_build/ref_callers/ (git-ignored; travels to the GPU box with the snapshot).  Test infrastructure:
tests/test_unmodified_reference_callers.py and __graft_entry__.build() call it; nothing of the reference is copied."""
from __future__ import annotations

import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
REF = Path("/root/reference")
OUT = ROOT / "tests" / "cpp" / "_build" / "ref_callers"
LIB_DIR = ROOT / "dynamicgraphrepresentationlearning_b200" / "lib"
MINI_GTEST = ROOT / "tests" / "cpp" / "mini_gtest"

DRIVERS = ["throughput-latency", "memory-throughput-latency", "memory-footprint", "vertex-classification"]      # experiments/src/*.cpp
TESTS = {"ref_test_wharfmh": "wharfmh.cpp", "ref_test_pairings": "pairings.cpp"}                                  # tests/*.cpp (googletest files)


def reference_present() -> bool:
    return (REF / "experiments" / "src" / "throughput-latency.cpp").exists()


def _compile(sources: list[Path], exe: Path, extra: list[str]) -> None:
    deps = sources + list((ROOT / "include").rglob("*.h*")) + [LIB_DIR / "libwharfmh_b200.so", Path(__file__)]
    if exe.exists() and all(exe.stat().st_mtime >= d.stat().st_mtime for d in deps if d.exists()):
        return
    cmd = ["g++", "-std=c++17", "-O1", "-w", *extra, f"-I{ROOT / 'include' / 'compat'}", f"-I{ROOT / 'include'}", *map(str, sources), "-o", str(exe),
           f"-L{LIB_DIR}", "-lwharfmh_b200", "-Wl,-rpath,$ORIGIN/../../../../dynamicgraphrepresentationlearning_b200/lib"]
    subprocess.run(cmd, check=True)


# the reference's own graph/wharfmh.h and the C ABI in ONE translation unit (tests/cpp/intree_stub.cpp),
# i.e. the unmodified reference engine and the B200 engine in one binary.  Flags of oracle/Makefile (-O2).
REFERENCE_FLAGS = ["-DEDGELONG", "-mcx16", "-std=c++17", "-march=x86-64-v2", "-Wno-subobject-linkage", "-w", "-DUSEMALLOC", "-DHOMEGROWN", "-pthread", "-O2"]


def build_intree_stub() -> Path:
    exe = OUT / "intree_stub"
    src = ROOT / "tests" / "cpp" / "intree_stub.cpp"
    deps = [src, ROOT / "include" / "wharfmh_b200.h", LIB_DIR / "libwharfmh_b200.so", Path(__file__)]
    if exe.exists() and all(exe.stat().st_mtime >= d.stat().st_mtime for d in deps if d.exists()):
        return exe
    OUT.mkdir(parents=True, exist_ok=True)
    inc = [f"-I{REF / d}" for d in ("libs/libcuckoo", "libs/compressed_trees", "graph", "walks", "utils", "config")]
    subprocess.run(["g++", *REFERENCE_FLAGS, *inc, f"-I{ROOT / 'include'}", str(src), "-o", str(exe), f"-L{LIB_DIR}", "-lwharfmh_b200",
                    "-Wl,-rpath,$ORIGIN/../../../../dynamicgraphrepresentationlearning_b200/lib", "-ldl"], check=True)
    return exe


def build_all() -> list[Path]:
    if not reference_present():
        raise FileNotFoundError("/root/reference is not on this machine")
    OUT.mkdir(parents=True, exist_ok=True)
    built = []
    for name in DRIVERS:
        exe = OUT / name
        _compile([REF / "experiments" / "src" / f"{name}.cpp"], exe, [])
        built.append(exe)
    for name, src in TESTS.items():
        exe = OUT / name
        _compile([REF / "tests" / src, MINI_GTEST / "main.cpp"], exe, [f"-I{MINI_GTEST}"])
        built.append(exe)
    built.append(build_intree_stub())
    return built


if __name__ == "__main__":
    for path in build_all():
        print(path)
