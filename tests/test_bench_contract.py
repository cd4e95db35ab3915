"""bench.py's host-side pieces, without a GPU: the reference arm's JSON line (the contract keys the driver reads), the
algorithmic-bytes model of a node2vec MH step and the clock sampler's rank gating."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def bench_workloads():
    import bench
    return bench.WORKLOADS


def test_default_run_puts_the_north_star_configuration_on_top():
    """`python bench.py --gpus N --steps K --warmup W` (what the driver runs) must headline RMAT scale-25 DeepWalk and carry
    the scale-20 node2vec MH block and the scale-14 pair of the reference arm."""
    import bench
    src = (ROOT / "bench.py").read_text()
    assert 'ap.add_argument("--workload", default="all"' in src
    assert 'run_workload(ctx, "dw25"' in src and '"n2v20": run_workload(ctx, "n2v20"' in src and '"dw14": run_workload(ctx, "dw14"' in src
    assert bench.WORKLOADS["dw25"]["n"] == 1 << 25 and bench.WORKLOADS["dw25"]["det"] and bench.WORKLOADS["dw25"]["model"] == "deepwalk"
    assert bench.REFERENCE_WORKLOAD == "dw14" and bench.WORKLOADS["dw14"]["edges_half"] == 8 * bench.WORKLOADS["dw14"]["n"]


def test_reference_arm_prints_one_contract_line(wo):
    if not wo.ref_available():
        pytest.skip("oracle/_ref (the unmodified reference) is not built on this machine")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "walk-steps/sec (init walks)" and d["unit"] == "walk-steps/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] >= 3          # timing rules: W >= 3
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    # the CPU arm runs BASELINE configs[0] in full: the size both arms share (the b200 line carries it as blocks.dw14)
    assert d["config"]["workload"] == bench_workloads()["dw14"]["desc"] and d["config"]["deterministic_mode"] is True
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert "the whole workload" in cb["sample"] and "scale-14" in cb["sample"] and "inverted index" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_mh_bytes_per_step_model():
    import torch
    import bench
    # degrees 1, 8, 64, 4096 -> S = 1, 1, 4, 10 sectors per binary search (S(d) = max(1, ceil(log2(d / 8)) + 1))
    offsets = torch.tensor([0, 1, 9, 73, 73 + 4096], dtype=torch.int64)
    for vertex, sectors in ((0, 1), (1, 1), (2, 4), (3, 10)):
        walks = torch.tensor([[vertex], [0]], dtype=torch.int64)         # position-major sample: prev at position 0
        assert bench.mh_bytes_per_step(torch, offsets, walks) == 32.0 * (4 + 2 * sectors) + 4.0
    stopped = torch.tensor([[2, 0xFFFFFFFF], [0, 0]], dtype=torch.int64)  # a stopped walker contributes no step
    assert bench.mh_bytes_per_step(torch, offsets, stopped) == 32.0 * (4 + 2 * 4) + 4.0
    assert bench.BYTES_PER_DET_STEP == 68.0


def test_clock_sampler_is_silent_on_other_ranks():
    import bench
    with bench.ClockSampler(0, enabled=False) as clocks:
        pass
    s = clocks.summary()
    assert s["sm_mhz"] is None and s["reasons"] == [] and "not sampled" in s["source"]
