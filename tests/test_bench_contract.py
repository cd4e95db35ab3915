"""bench.py's host-side pieces, without a GPU: the reference arm's JSON line (the contract keys the driver reads), the
algorithmic-bytes model of a node2vec MH step and the clock sampler's rank gating."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def test_reference_arm_prints_one_contract_line(wo):
    if not wo.ref_available():
        pytest.skip("oracle/_ref (the unmodified reference) is not built on this machine")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-sample-scale", "12"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "walk-steps/sec (init walks)" and d["unit"] == "walk-steps/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] >= 3          # timing rules: W >= 3
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert "workload" in d["config"] and "node2vec MH" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "scale-12" in cb["sample"]
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
