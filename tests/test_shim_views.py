"""Snapshot views (flatten_vertex_tree, map_vertices) and the walks.txt writer of the C++ re-creation, read the way the
reference's tests read the originals (tests/wharfmh.cpp:65-99,108-139; vertex-classification.cpp:167-178).  Compiles on
CPU; runs on the GPU box against the six-vertex corpora of the unmodified reference (tests/golden/golden.json)."""
import json
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
LIB_DIR = ROOT / "dynamicgraphrepresentationlearning_b200" / "lib"
BUILD = ROOT / "tests" / "cpp" / "_build"


def build_driver() -> Path:
    BUILD.mkdir(parents=True, exist_ok=True)
    exe = BUILD / "shim_views_driver"
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", f"-I{ROOT / 'include'}", str(ROOT / "tests" / "cpp" / "shim_views_driver.cpp"),
                    "-o", str(exe), f"-L{LIB_DIR}", "-lwharfmh_b200", f"-Wl,-rpath,{LIB_DIR}"], check=True)
    return exe


def test_shim_views_compile_and_fail_loudly_without_gpu(tmp_path):
    exe = build_driver()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_shim_views_on_gpu")
    r = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_shim_views_on_gpu(tmp_path):
    exe = build_driver()
    out = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True, check=True).stdout.splitlines()
    six = json.loads((ROOT / "tests" / "golden" / "golden.json").read_text())["six"]
    twos = sum(w.count(2) for w in six["walks"][0])
    assert "fresh tree 6 vertices 0 index entries 0 roots degree(2) 5 samplers 0" in out
    assert f"vertex tree 6 vertices 96 index entries, vertex 2: degree 5 first 0 last 5 index {twos} root 1 samplers 0 mapped 6" in out
    assert any(l.startswith("walks.txt ") and l.endswith(" 12 lines same 1") for l in out)       # all 12 walks affected (golden)
    assert (tmp_path / "walks.txt").read_text() == "".join("".join(f"{v} " for v in w) + "\n" for w in six["walks"][1])
    assert "after destroy_index 6 vertices 0 index entries degree(0) 3" in out
    assert "destroyed 0 vertices tree 0" in out
