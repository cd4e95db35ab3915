#!/usr/bin/env python
"""Headline benchmark of the WharfMH hot path on B200 (contract: see the round brief, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A *step* is one full pass of initial walk generation (graph/wharfmh.h:250) over the workload's graph:
10 walks per vertex of length 80.  `value` = walk-steps/s with the graph resident in HBM, timed with CUDA
events on the engine's stream; `e2e` = the same through the C ABI with host buffers (graph H2D, generate,
corpus D2H inside the timed region).  At N > 1 the walks shard over the ranks (whole replicas first, the rest by
start vertex; graph replicated, no data-path collective while walking) and the same total work is timed: strong scaling.

Workloads (BASELINE.json configs):
    dw25   configs[3]: DeepWalk on RMAT scale-25 (33M V, 537M E) - the configuration the north star quotes its target
           on; fits one B200 (115 GB of 180 GB)                                           <- headline of the default run
    n2v20  configs[1]: node2vec MH (p=0.5, q=2) on RMAT scale-20 (1M V, 16M E) + edge-batch updates <- second block
    dw14   configs[0]: DeepWalk on RMAT scale-14 (the reference's CPU-runnable case): the like-for-like pair of the
           reference arm (`--impl reference` runs exactly this, in full)                   <- third block
    lj     configs[2]: LiveJournal-shape RMAT (4.8M V, 69M E) node2vec MH (+ 10 insert batches of 100K)
    n2v23  configs[4]: RMAT scale-23 node2vec MH (batch-size sweep with --batch-sweep)
`--workload all` (default) prints ONE line: dw25 on top, `blocks.n2v20` and `blocks.dw14` inside; naming one workload
makes it the line's headline.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    #        scale/vertices, directed-edge target, model, deterministic
    "n2v20": dict(n=1 << 20, vertices=(1 << 20) + 1, edges_half=8 << 20, model="node2vec", det=False, scale=20, desc="node2vec MH (p=0.5,q=2) 10x80, RMAT scale-20 (1M V, 16M E)"),
    "dw14": dict(n=1 << 14, vertices=(1 << 14) + 1, edges_half=128 << 10, model="deepwalk", det=True, scale=14, desc="DeepWalk 10x80, RMAT scale-14 (16K V, 256K E)"),
    "dw25": dict(n=1 << 25, vertices=(1 << 25) + 1, edges_half=268435456, model="deepwalk", det=True, scale=25, desc="DeepWalk 10x80, RMAT scale-25 (33M V, 537M E)"),
    "lj": dict(n=4847571, vertices=4847571, edges_half=34500000, model="node2vec", det=False, scale=None, desc="node2vec MH 10x80, LiveJournal-shape RMAT (4.8M V, 69M E)"),
    "n2v23": dict(n=1 << 23, vertices=(1 << 23) + 1, edges_half=67108864, model="node2vec", det=False, scale=23, desc="node2vec MH (p=0.5,q=2) 10x80, RMAT scale-23 (8.4M V, 134M E)"),
}
WPV, LEN, P, Q = 10, 80, 0.5, 2.0
GRAPH_SEED = 42
MH_SEED = 1
BYTES_PER_DET_STEP = 68.0      # 32 B row descriptor sector + 32 B neighbour sector + 4 B coalesced store
REFERENCE_WORKLOAD = "dw14"    # BASELINE.json configs[0]: "run by the reference on CPU" - the size both arms share
MASK64 = (1 << 64) - 1


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ----------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi during the timed region
# ----------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled every 20 ms from a thread of RANK 0 only (the
    timed region of the default run is well under a second, one nvidia-smi process launch would not fit into it);
    nvidia-smi when NVML is unavailable.  Other ranks do not poll: NVML queries take driver-wide locks, and eight
    processes polling every few milliseconds showed up as ~1 ms of launch gaps inside a 4 ms multi-launch pass."""
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index: int, enabled: bool = True):
        self.gpu = gpu_index
        self.enabled = enabled
        self.sm: list[int] = []
        self.max_sm = None
        self.reasons: set[str] = set()
        self.source = "nvml"
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        if not enabled:
            self.source = "not sampled on this rank"
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            physical = int(visible.split(",")[gpu_index]) if visible and all(x.strip().isdigit() for x in visible.split(",")) else gpu_index
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(physical)
            self.max_sm = int(pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:
            self.source = "nvidia-smi"

    def _poll_nvml(self):
        nv = self._nvml
        masks = [(nv.nvmlClocksEventReasonHwSlowdown, "hw_slowdown"), (nv.nvmlClocksEventReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                 (nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown"), (nv.nvmlClocksEventReasonSwPowerCap, "sw_power_cap")]
        while not self._stop.is_set():
            try:
                self.sm.append(int(nv.nvmlDeviceGetClockInfo(self._handle, nv.NVML_CLOCK_SM)))
                bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._handle))
                self.reasons.update(name for mask, name in masks if bits & mask)
            except Exception:
                pass
            self._stop.wait(0.02)

    def _poll_smi(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    f = [x.strip() for x in out.splitlines()[0].split(",")]
                    if f[0].isdigit():
                        self.sm.append(int(f[0]))
                    if len(f) > 1 and f[1].isdigit():
                        self.max_sm = int(f[1])
                    self.reasons.update(self.NAMES[i] for i in range(4) if len(f) > 2 + i and f[2 + i].lower().startswith("active"))
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        if self.enabled:
            self._thread = threading.Thread(target=self._poll_nvml if self._nvml else self._poll_smi, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=6)

    def summary(self) -> dict:
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_sm, "reasons": sorted(self.reasons),
                "samples": len(self.sm), "source": self.source}


# ----------------------------------------------------------------------------------------------------
# synthetic input: the reference's own generator (utils/utility.h:55-146), restated on the device
# ----------------------------------------------------------------------------------------------------
def make_graph_on_device(wharf, torch, spec, device_index):
    """generate_batch_of_edges(E/2, vertices, seed 42, self_loops=false, directed=false) -> symmetric CSR."""
    lib = wharf._capi.load()
    n = spec["n"]          # vertices = 2^s + 1 makes the generator emit ids < 2^s (utility.h:78); SURVEY.md 8d
    total = 2 * spec["edges_half"]
    dev = torch.device(f"cuda:{device_index}")
    pairs = torch.empty(2 * total, dtype=torch.int32, device=dev)
    k = C.c_uint64()
    wharf._capi.check(lib.wmh_generate_batch_of_edges_device(device_index, spec["edges_half"], spec["vertices"], GRAPH_SEED, 0, 0,
                                                             0.5, 0.2, 0.1, pairs.data_ptr(), C.byref(k)))
    m = int(k.value)
    offsets = torch.empty(n + 1, dtype=torch.int64, device=dev)
    edges = torch.empty(max(m, 1), dtype=torch.int32, device=dev)
    wharf._capi.check(lib.wmh_pairs_to_csr_device(device_index, n, pairs.data_ptr(), m, offsets.data_ptr(), edges.data_ptr()))
    del pairs
    torch.cuda.empty_cache()
    wharf.trim_memory(device_index)      # the generator's sort buffers go back to the driver
    return n, m, offsets, edges


def make_batch_on_device(wharf, torch, edges_number, n, seed, device_index):
    lib = wharf._capi.load()
    out = torch.empty(4 * edges_number, dtype=torch.int32, device=f"cuda:{device_index}")
    k = C.c_uint64()
    wharf._capi.check(lib.wmh_generate_batch_of_edges_device(device_index, edges_number, n, seed, 0, 0, 0.5, 0.2, 0.1, out.data_ptr(), C.byref(k)))
    return out[: 2 * int(k.value)].contiguous(), int(k.value)


def mh_bytes_per_step(torch, offsets, walks_sample):
    """SURVEY.md 8d node2vec-MH step: 32 B x (desc[cur] + candidate + desc[prev] + sampler state + 2 x S(deg(prev))) + 4 B,
    S(d) = max(1, ceil(log2(d/8)) + 1) sectors per binary search; averaged over a sample of the generated trace."""
    deg = (offsets[1:] - offsets[:-1]).to(torch.float64)
    prev = walks_sample[:-1].reshape(-1).to(torch.int64)           # vertex at position p is prev of the step at p+1
    prev = prev[prev != 0xFFFFFFFF]
    d = deg[prev].clamp(min=1)
    s = torch.clamp(torch.ceil(torch.log2(d / 8)) + 1, min=1)
    return float(32.0 * (4 + 2 * s.mean().item()) + 4.0)


# ----------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the unmodified reference (oracle/_ref) on the host cores
# ----------------------------------------------------------------------------------------------------
def cpu_reference_run(spec, trials: int, sample_scale: int | None):
    """Runs the reference's generate_initial_random_walks `trials` times on the workload itself when it is CPU-sized
    (sample_scale None: BASELINE configs[0]) or on a bounded sample of it (same generator and parameters, smaller
    scale).  The timed call builds the walks AND the inverted index (graph/wharfmh.h:250-356).
    Returns (per-trial seconds, steps per trial, cores, kind, description)."""
    from oracle import wharf_oracle as wo
    full = sample_scale is None
    scale = spec["scale"] if full else sample_scale
    n = 1 << scale
    off, adj = wo.rmat_graph(scale, 8 * n, GRAPH_SEED)
    cores = os.cpu_count() or 1
    what = "the whole workload" if full else f"bounded sample of {spec['desc']}"
    sample = f"{what}: RMAT scale-{scale} ({n} V, {adj.shape[0]} E, generator seed {GRAPH_SEED}), full 10x80 corpus incl. inverted index, {trials} trial(s)"
    if wo.ref_available():
        with tempfile.TemporaryDirectory() as tmp:
            g = Path(tmp) / "g.graph"
            wo.write_graph(g, n, off, adj)
            res = wo.ref_run(["run", g, Path(tmp) / "out", f"model={spec['model']}", f"p={P}", f"q={Q}", f"det={int(spec['det'])}",
                              f"wpv={WPV}", f"len={LEN}", "dump=0", f"trials={trials}"], threads=cores, timeout=3000)
        init = res[0]
        return init["init_times"], init["steps"], cores, "reference", sample
    log("oracle/_ref/wharf_ref absent: timing the plain-C port instead")
    model = wo.NODE2VEC if spec["model"] == "node2vec" else wo.DEEPWALK
    o = wo.Oracle(n, off, adj, walks_per_vertex=WPV, walk_length=LEN, model=model, p=P, q=Q, deterministic=spec["det"])
    times = []
    for _ in range(trials):
        t0 = time.perf_counter()
        steps = o.generate()
        times.append(time.perf_counter() - t0)
    return times, steps, wo.lib().wo_threads(), "port", sample


def reference_spec(args):
    """The workload the CPU arm runs: the named one when the reference can hold it (dw14), else configs[0]."""
    name = args.workload if args.workload in ("dw14",) else REFERENCE_WORKLOAD
    return name, WORKLOADS[name]


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name, spec = reference_spec(args)
    trials = args.warmup + args.steps
    times, steps, cores, kind, sample = cpu_reference_run(spec, trials, None)
    timed = times[args.warmup:]
    value = steps * len(timed) / sum(timed)
    line = {
        "impl": "reference", "metric": "walk-steps/sec (init walks)", "value": value, "unit": "walk-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(timed) / len(timed), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": spec["desc"], "walks_per_vertex": WPV, "walk_length": LEN, "deterministic_mode": spec["det"], "sample": sample,
                   "note": "BASELINE.json configs[0], run in full: the size both arms share.  The b200 arm's line carries the same workload as "
                           "`blocks.dw14` (value_with_index is the like-for-like figure: the reference's call builds its inverted index too); its "
                           "headline dw25 cannot run here at all (32-bit walk keys wrap beyond 5.37 M vertices, walks/inverted_index.h:14,76)"},
        "cpu_baseline": {"value": value, "unit": "walk-steps/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "walk-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------------------------------
class Ctx:
    """Rank plumbing shared by the workload blocks."""

    def __init__(self, args, torch, dist, wharf, np):
        self.args, self.torch, self.dist, self.wharf, self.np = args, torch, dist, wharf, np
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dev = torch.device(f"cuda:{self.local_rank}")
        self.flush = None
        peaks_path = ROOT / "MEASURED_PEAKS.json"
        peaks = json.loads(peaks_path.read_text()) if peaks_path.exists() else {}
        self.peak_gbs, self.peak_kind = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
        tpath = ROOT / "profiles" / "roofline_traffic.json"
        self.traffic = json.loads(tpath.read_text()) if tpath.exists() else {}
        fpath = ROOT / "tests" / "golden" / "bench_fingerprints.json"
        self.golden_fp = json.loads(fpath.read_text()) if fpath.exists() else {}

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x: int) -> int:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return int(t.item())

    def wrapping_sum_over_ranks(self, x: int) -> int:
        """Sum modulo 2^64 of one u64 per rank (corpus fingerprints are wrapping sums of per-walk hashes)."""
        if self.world == 1:
            return x & MASK64
        parts = self.torch.tensor([x & 0xFFFFFFFF, x >> 32], dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(parts, op=self.dist.ReduceOp.SUM)
        lo, hi = (int(v) for v in parts.tolist())
        return (lo + (hi << 32)) & MASK64

    def flush_l2(self):
        if self.flush is None:
            self.flush = self.torch.empty(256 << 20, dtype=self.torch.uint8, device=self.dev)       # > 126 MB L2
        self.flush.fill_(1)


def timed_passes(ctx, engine, steps, warmup, peers):
    """`warmup` untimed + `steps` timed generate passes; device time (CUDA events on the engine stream) per pass."""
    torch = ctx.torch

    def one_step():
        ctx.flush_l2()                                                   # L2 flush between timed iterations
        torch.cuda.synchronize()
        if peers:
            ctx.barrier()                                                # ranks enter the pass together: the peer exchange meets at a flag
                                                                         # barrier ON the device, so host-side skew would be timed as waiting
        s = engine.generate_initial_random_walks()                       # device-timed inside (CUDA events on the engine stream)
        return s, engine.last_walk_kernel_ms(), engine.last_index_ms()

    for _ in range(warmup):
        one_step()
    launches0 = engine.kernel_launches()
    ctx.barrier()
    wall0 = time.perf_counter()
    kernel_ms, index_ms, local_steps = [], [], 0
    for _ in range(steps):
        s, ms, ims = one_step()
        local_steps += s
        kernel_ms.append(ms)
        index_ms.append(ims)
    ctx.barrier()
    wall = time.perf_counter() - wall0
    return {"kernel_ms": kernel_ms, "index_ms": index_ms, "local_steps": local_steps, "wall": wall, "launches": engine.kernel_launches() - launches0}


def roofline_for(ctx, name, spec, n, m, d_offsets, engine, local_steps_per_pass, kernel_avg_ms, coalesced):
    """Roofline of the dominant kernel(s) of one generate pass on rank 0.  `achieved` = DRAM bytes the pass moved (one
    ncu capture of this workload and world size, profiles/roofline_traffic.json) / its device time measured live;
    the SURVEY 8d algorithmic figure rides beside it (it is not a bound for a schedule that fetches a row once for many
    walkers, nor for a kernel half of whose sectors hit the L2)."""
    torch, np = ctx.torch, ctx.np
    if spec["det"] or spec["model"] != "node2vec":
        bytes_per_step = BYTES_PER_DET_STEP
    else:
        sample = torch.from_numpy(engine.walks(0, min(20000, engine.number_of_local_walks())).astype(np.int64)).to(ctx.dev).t().contiguous()
        bytes_per_step = mh_bytes_per_step(torch, d_offsets, sample)
    if spec["det"]:
        kernel_name = ("coalesced_{advance,tile_counts,expand,relabel,emit}_kernel segments (kernels_coalesced.cu), one generate pass replayed as a CUDA graph"
                       if coalesced else "walk_deterministic_ilp_kernel<1,256,0>")
    else:
        kernel_name = ("walk_n2v_kernel<false> (+ n2v_share_init_kernel: bulk sampler init)" if spec["model"] == "node2vec" else "walk_deepwalk_mh_kernel<false>")
    algorithmic_gbs = local_steps_per_pass * bytes_per_step / (kernel_avg_ms * 1e-3) / 1e9
    t = ctx.traffic.get(f"{name}:{ctx.world}")
    out = {"bound": "hbm", "peak": ctx.peak_gbs, "unit": "GB/s", "kernel": kernel_name, "kernel_ms": kernel_avg_ms,
           "steps_per_launch": local_steps_per_pass, "peak_source": ctx.peak_kind,
           "algorithmic": {"bytes_per_step": bytes_per_step, "achieved": algorithmic_gbs, "frac": algorithmic_gbs / ctx.peak_gbs,
                           "note": "SURVEY 8d bytes per walker step x steps / time: an upper estimate of the traffic of a one-walker-per-thread "
                                   "kernel whose loads all miss; shown for comparison, not as this pass's bound"}}
    if isinstance(t, dict):
        traffic = float(t["read"] + t["write"])
        out.update({"achieved": traffic / (kernel_avg_ms * 1e-3) / 1e9, "traffic": traffic, "traffic_read": t["read"], "traffic_write": t["write"],
                    "traffic_source": t.get("source", "profiles/roofline_traffic.json"),
                    "definition": "achieved = measured DRAM bytes of one pass (ncu dram__bytes_read.sum + dram__bytes_write.sum over its kernels, this "
                                  "workload and world size) / device time of the pass measured here; frac = achieved / peak"})
        out["frac"] = out["achieved"] / ctx.peak_gbs
        out["bytes_per_step_measured"] = traffic / local_steps_per_pass
    else:
        frac = algorithmic_gbs / ctx.peak_gbs
        if frac <= 1.0 and n * 8 + m * 4 > (126 << 20):      # an L2-resident graph is no HBM workload at any rate
            out.update({"achieved": algorithmic_gbs, "traffic": None, "frac": frac,
                        "definition": "no ncu capture for this workload / world size: achieved falls back to the SURVEY 8d algorithmic bytes"})
        else:       # the model does not describe what ran (L2-resident graph, or rows fetched once for many walkers): no HBM fraction is claimed
            out.update({"achieved": None, "traffic": None, "frac": None,
                        "definition": "no ncu capture for this workload / world size, and the SURVEY 8d algorithmic bytes exceed what HBM can move in the "
                                      "measured time (cached rows): no HBM fraction claimed; see `algorithmic`"})
    if spec["det"]:
        corpus_bytes = float(engine.number_of_local_walks()) * LEN * 4
        floor_ms = corpus_bytes / (ctx.peak_gbs * 1e9) * 1e3
        out["store_floor"] = {"corpus_bytes": corpus_bytes, "floor_ms": floor_ms, "frac_of_floor": floor_ms / kernel_avg_ms,
                              "note": "every corpus cell is computed and stored once: corpus bytes / peak is the floor of any schedule"}
    return out


def pinned_empty(torch, np, shape, dtype):
    """Page-locked host array of exactly the size asked for (cudaHostRegister over a numpy allocation): torch's pinned
    allocator rounds every block up to a power of two, which turns the 107 GB corpus buffer into 128 GB."""
    a = np.empty(shape, dtype=dtype)
    if a.nbytes:
        err = torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)
        if int(err) != 0:
            raise RuntimeError(f"cudaHostRegister of {a.nbytes} bytes failed: {err}")
    return a


def pinned_release(torch, a):
    if a.nbytes:
        torch.cuda.cudart().cudaHostUnregister(a.ctypes.data)


def run_e2e(ctx, spec, cfg, n, m, d_offsets, d_edges, iters):
    """Through the C ABI with HOST buffers: wmh_create (graph H2D) + wmh_generate_initial_random_walks_to_host (the whole corpus
    D2H into pinned memory while it is generated) + wmh_destroy, all inside the timed region."""
    torch, np, wharf = ctx.torch, ctx.np, ctx.wharf
    probe = wharf.WharfMH(n, m, d_offsets.data_ptr(), d_edges.data_ptr(), config=cfg, device_arrays=True)
    w_count = probe.number_of_local_walks()
    probe.destroy()
    corpus_bytes = w_count * LEN * 4
    graph_bytes = n * 8 + m * 4
    # every rank of the box pins its share of the corpus: refuse when the box's host memory cannot hold it comfortably
    ctx.barrier()
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = None
    ctx.barrier()
    if avail is not None and ctx.world * (corpus_bytes + graph_bytes) > 0.75 * avail:
        return {"value": None, "unit": "walk-steps/s", "h2d_bytes_per_step": int(graph_bytes), "d2h_bytes_per_step": int(corpus_bytes),
                "unavailable": f"host memory: {ctx.world} rank(s) x {(corpus_bytes + graph_bytes) >> 30} GiB of pinned buffers, {avail >> 30} GiB available on the box"}
    t_pin = time.perf_counter()
    h_off_np, h_adj_np = pinned_empty(torch, np, n, np.uint64), pinned_empty(torch, np, max(m, 1), np.uint32)
    h_walks_np = pinned_empty(torch, np, (w_count, LEN), np.uint32)
    torch.from_numpy(h_off_np.view(np.int64)).copy_(d_offsets[:n])
    torch.from_numpy(h_adj_np.view(np.int32)).copy_(d_edges[: max(m, 1)])
    torch.cuda.synchronize()
    pin_s = time.perf_counter() - t_pin
    phases = []

    def e2e_step():
        t = [time.perf_counter()]
        eng = wharf.WharfMH(n, m, h_off_np, h_adj_np, config=cfg)                                             # H2D
        t.append(time.perf_counter())
        s = eng.generate_initial_random_walks_to_host(h_walks_np)                                             # generate + D2H, overlapped
        t.append(time.perf_counter())
        eng.destroy()
        t.append(time.perf_counter())
        phases.append([round(1e3 * (b - a), 2) for a, b in zip(t, t[1:])])
        return s

    e2e_step()
    ctx.barrier()
    t0 = time.perf_counter()
    e_steps = 0
    for _ in range(iters):
        e_steps += e2e_step()
    ctx.barrier()
    e2e_s = ctx.max_over_ranks(time.perf_counter() - t0)
    # the host copy is the product: its fingerprint must equal the device corpus's (computed on the host for a slice)
    out = {"value": ctx.sum_over_ranks(e_steps) / e2e_s, "unit": "walk-steps/s",
           "h2d_bytes_per_step": int(n * 8 + m * 4), "d2h_bytes_per_step": int(corpus_bytes),
           "iters": iters, "ms_per_step": 1e3 * e2e_s / iters, "pin_host_buffers_s": round(pin_s, 2),
           "phases_ms": {"order": ["create_h2d", "generate_to_host", "destroy"], "warm": phases[0], "timed": phases[1:]},
           "path": "wmh_create(host CSR) + wmh_generate_initial_random_walks_to_host(pinned host: whole corpus D2H overlapped with the walk kernels) + wmh_destroy, per rank"}
    for a in (h_walks_np, h_off_np, h_adj_np):
        pinned_release(torch, a)
    del h_walks_np, h_off_np, h_adj_np
    return out


def run_updates(ctx, engine, n, sizes, insert_only, search_name):
    """Edge-batch update latency through the public call: at N = 1 the host C-ABI call (H2D of the batch, merge, search,
    re-walk); at N > 1 ShardedWharfMH.insert/delete_edges_batch - rank 0 owns the batch, C1 NCCL broadcast, the local
    update on every rank, C2 NCCL all-gather of the affected walk ids - all inside the measured latency."""
    torch, np, wharf = ctx.torch, ctx.np, ctx.wharf
    sharded = None
    if ctx.world > 1:
        from dynamicgraphrepresentationlearning_b200.distributed import ShardedWharfMH
        sharded = ShardedWharfMH.from_engine(engine)
        sharded.warm_collectives(engine.number_of_local_walks())     # NCCL channel and buffer set-up is not batch latency
    names = {wharf.SEARCH_INDEX: "index", wharf.SEARCH_SCAN: "scan", 0: "none"}
    rows = []
    for b, size in enumerate(sizes):
        d_batch, k = make_batch_on_device(wharf, torch, size, n, b, ctx.local_rank)
        h_batch = torch.empty(2 * k, dtype=torch.int32, pin_memory=True)
        h_batch.copy_(d_batch)
        torch.cuda.synchronize()
        hb = h_batch.numpy().view(np.uint32).reshape(-1, 2)
        row = {"batch_edges": k}
        for kind in (("insert",) if insert_only else ("insert", "delete")):
            engine.reset_timers()
            ctx.barrier()
            t0 = time.perf_counter()
            if sharded is None:
                fn = engine._lib.wmh_insert_edges_batch if kind == "insert" else engine._lib.wmh_delete_edges_batch
                na = C.c_uint64()
                wharf._capi.check(fn(engine._h, k, C.c_void_p(hb.ctypes.data), 1, 0, (1 << 64) - 1, 1, C.byref(na)))
                affected_total = int(na.value)
                timing = None
            else:
                fn = sharded.insert_edges_batch if kind == "insert" else sharded.delete_edges_batch
                ids = fn(hb if ctx.rank == 0 else None, src=0, return_device=True)
                affected_total = int(ids.shape[0])
                timing = dict(sharded.last_timing)
            ctx.barrier()
            dt = ctx.max_over_ranks(time.perf_counter() - t0)
            gi = 0 if kind == "insert" else 2
            row[kind] = {"latency_ms": 1e3 * dt, "graph_update_ms": 1e3 * ctx.max_over_ranks(engine.timer(gi)),
                         "walk_update_ms": 1e3 * ctx.max_over_ranks(engine.timer(gi + 1)),
                         "affected_walks": affected_total if sharded is not None else ctx.sum_over_ranks(affected_total),
                         "rewalk_steps": ctx.sum_over_ranks(engine.last_steps()), "rewalk_kernel_ms": ctx.max_over_ranks(engine.last_walk_kernel_ms()),
                         "index_ms": ctx.max_over_ranks(engine.last_index_ms()), "search": names.get(engine.last_search_mode(), "?")}
            if timing:
                row[kind].update({"bcast_ms": ctx.max_over_ranks(timing["bcast_ms"]), "local_update_ms": ctx.max_over_ranks(timing["update_ms"]),
                                  "allgather_ms": ctx.max_over_ranks(timing["allgather_ms"])})
        rows.append(row)
    ins = [r["insert"]["latency_ms"] for r in rows]
    dele = [r["delete"]["latency_ms"] for r in rows if "delete" in r] or [float("nan")]
    path = ("wmh_insert/delete_edges_batch host call (H2D batch, merge, affected search, re-walk)" if sharded is None else
            "ShardedWharfMH.insert/delete_edges_batch: C1 NCCL broadcast of rank 0's batch + wmh_*_edges_batch_device on every rank + C2 NCCL "
            "all-gather (counts, padded ids) and device sort of the affected global walk ids")
    return {"search_requested": search_name, "batches": rows, "insert_latency_ms_median": statistics.median(ins),
            "delete_latency_ms_median": statistics.median(dele), "host_batch": True, "collectives_in_latency": sharded is not None,
            "note": "latency = " + path + ", max over ranks; `search` = how the affected walks were found, as taken"}


def index_vs_scan(ctx, engine, n, sizes):
    """The same batch on the same state under WMH_SEARCH_INDEX and WMH_SEARCH_SCAN (device-resident batch, local update
    only): what the AUTO rule chooses between.  INDEX pays the search in the per-vertex segments plus keeping the index
    current; SCAN streams the walk matrix once."""
    torch, wharf = ctx.torch, ctx.wharf
    out = []
    # one untimed insert + delete first: the first batches after a layout switch pay for pool growth and a compaction
    d_warm, k_warm = make_batch_on_device(wharf, torch, max(sizes), n, 999, ctx.local_rank)
    engine.insert_edges_batch_device(d_warm.data_ptr(), k_warm)
    engine.delete_edges_batch_device(d_warm.data_ptr(), k_warm)
    for size in sizes:
        d_batch, k = make_batch_on_device(wharf, torch, size, n, 1000 + size, ctx.local_rank)
        # one untimed cycle of this very batch: its delete also removes the batch's edges that were in the graph before, so
        # only from the second cycle on do both modes start from the same graph (and report the same affected walks)
        engine.insert_edges_batch_device(d_batch.data_ptr(), k)
        engine.delete_edges_batch_device(d_batch.data_ptr(), k)
        row = {"batch_edges": k}
        for mode_name, mode in (("index", wharf.SEARCH_INDEX), ("scan", wharf.SEARCH_SCAN)):
            engine.destroy_index()
            engine.set_config(search_mode=mode)
            try:
                engine.generate_initial_random_walks()                   # INDEX: builds the index inside
            except wharf.WharfError as ex:
                row[mode_name] = {"unavailable": str(ex)}
                continue
            lat = {}
            for kind, fn in (("insert", engine.insert_edges_batch_device), ("delete", engine.delete_edges_batch_device)):
                ctx.barrier()
                t0 = time.perf_counter()
                fn(d_batch.data_ptr(), k)
                ctx.barrier()
                lat[kind] = {"latency_ms": 1e3 * ctx.max_over_ranks(time.perf_counter() - t0),
                             "affected_walks": ctx.sum_over_ranks(engine.number_of_affected_walks()),
                             "index_ms": ctx.max_over_ranks(engine.last_index_ms()), "rewalk_kernel_ms": ctx.max_over_ranks(engine.last_walk_kernel_ms()),
                             "search": {wharf.SEARCH_INDEX: "index", wharf.SEARCH_SCAN: "scan"}.get(engine.last_search_mode(), "none")}
            row[mode_name] = lat
        out.append(row)
    return out


def run_workload(ctx, name, *, steps, warmup, e2e_iters, updates, headline):
    """One workload block: generate passes (walk only, and with the inverted index built inside the call where it fits),
    corpus fingerprints, roofline, e2e, and for MH workloads the deterministic-mode figure and the edge-batch updates."""
    torch, dist, wharf, np, args = ctx.torch, ctx.dist, ctx.wharf, ctx.np, ctx.args
    spec = WORKLOADS[name]
    rank, world, local_rank = ctx.rank, ctx.world, ctx.local_rank
    t0 = time.perf_counter()
    n, m, d_offsets, d_edges = make_graph_on_device(wharf, torch, spec, local_rank)
    torch.cuda.synchronize()
    log(f"[rank {rank}] {name}: graph n={n} m={m} generated in {time.perf_counter() - t0:.2f}s")
    model = wharf.NODE2VEC if spec["model"] == "node2vec" else wharf.DEEPWALK
    cfg = wharf.Config(walks_per_vertex=WPV, walk_length=LEN, random_walk_model=model, paramP=P, paramQ=Q,
                       deterministic_mode=spec["det"], seed=MH_SEED, search_mode=wharf.SEARCH_SCAN, device=local_rank, rank=rank, world=world)
    engine = wharf.WharfMH(n, m, d_offsets.data_ptr(), d_edges.data_ptr(), config=cfg, device_arrays=True)
    peers = False
    mh_n2v = (not spec["det"]) and spec["model"] == "node2vec"
    if 1 < world <= 8 and mh_n2v and not args.no_peers:
        from dynamicgraphrepresentationlearning_b200.distributed import connect_peers
        peers = connect_peers(engine)                                    # exchange buffers mapped between the ranks (NVLink)
    coalesced = spec["det"] and n * 8 + m * 4 > (96 << 20)               # engine's GENERATE_AUTO rule (engine.cu use_coalesced)

    # ---- walk generation, corpus kept as the walk matrix (search_mode SCAN: no index is built) ----
    clock_ctx = ClockSampler(local_rank, enabled=rank == 0 and headline)
    with clock_ctx as clocks:
        run = timed_passes(ctx, engine, steps, warmup, peers)
    device_s = ctx.max_over_ranks(sum(run["kernel_ms"]) * 1e-3)
    total_steps = ctx.sum_over_ranks(run["local_steps"])
    value = total_steps / device_s
    kernel_avg_ms = sum(run["kernel_ms"]) / len(run["kernel_ms"])
    fp_key = "mh_fingerprint" if not spec["det"] else "det_fingerprint"
    parity = {fp_key: ctx.wrapping_sum_over_ranks(engine.corpus_fingerprint()), "steps_per_pass": total_steps // steps}
    block = {"workload": spec["desc"], "n": n, "m": m, "value": value, "unit": "walk-steps/s", "ms_per_step": 1e3 * device_s / steps,
             "steps": steps, "warmup": warmup, "gpu_launches": run["launches"], "wall_s_timed_region": run["wall"],
             "peer_exchanges": engine.peer_exchanges(), "deterministic_mode": spec["det"],
             "l2": "inputs larger than L2 (graph 2.4 GB, corpus 107 GB) and an explicit 256 MiB write between timed steps" if name == "dw25"
                   else "explicit 256 MiB write between timed steps"}
    roofline = roofline_for(ctx, name, spec, n, m, d_offsets, engine, run["local_steps"] / steps, kernel_avg_ms, coalesced)

    # ---- N > 1, MH: one untimed pass with every rank computing all sampler init values itself must give the same corpus ----
    if peers:
        engine.peer_detach()
        engine.generate_initial_random_walks()
        parity["mh_fingerprint_without_peers"] = ctx.wrapping_sum_over_ranks(engine.corpus_fingerprint())
        parity["peers_equal_no_peers"] = parity["mh_fingerprint_without_peers"] == parity[fp_key]
        from dynamicgraphrepresentationlearning_b200.distributed import connect_peers
        peers = connect_peers(engine)

    # ---- the reference-equivalent call: walks AND the inverted index built inside it (graph/wharfmh.h:312-354) ----
    cells = engine.number_of_local_walks() * LEN
    idx_steps = max(1, min(steps, 5))
    chunk_cells = min(cells, 1 << 31)                        # the index sorts the corpus in chunks of whole positions (32-bit cell ids)
    index_need = 4 * cells + 16 * chunk_cells + 8 * (n + 1) * -(-cells // max(chunk_cells, 1))
    fits = index_need < 0.7 * torch.cuda.mem_get_info(ctx.dev)[0]
    if ctx.world > 1:
        fits = ctx.sum_over_ranks(0 if fits else 1) == 0
    if fits:
        engine.set_config(search_mode=wharf.SEARCH_INDEX)
        run_i = timed_passes(ctx, engine, idx_steps, 1, peers)
        with_index_s = ctx.max_over_ranks((sum(run_i["kernel_ms"]) + sum(run_i["index_ms"])) * 1e-3)
        block["value_with_index"] = ctx.sum_over_ranks(run_i["local_steps"]) / with_index_s
        block["index_build_ms"] = ctx.max_over_ranks(sum(run_i["index_ms"]) / idx_steps)
        block["with_index_ms_per_step"] = 1e3 * with_index_s / idx_steps
        block["index_note"] = ("search_mode INDEX: the per-vertex inverted index (one radix sort of the corpus cells by vertex) is built inside the timed "
                               "call, as the reference's generate_initial_random_walks does; `value` is the same call with search_mode SCAN "
                               "(the walk matrix is the corpus, affected walks are found by scanning it)")
        engine.destroy_index()
        engine.set_config(search_mode=wharf.SEARCH_SCAN)
    else:
        block["value_with_index"] = None
        block["index_note"] = (f"no inverted index at this size on {world} GPU(s): {cells / 1e9:.1f} G cells per rank need {index_need / 1e9:.0f} GB "
                               "(4 B per cell + the sort buffers of one 2^31-cell chunk) beside the corpus.  The corpus is the walk matrix and "
                               "affected walks are found by scanning it (search_mode SCAN), which needs no index")

    # ---- deterministic schedules side by side on the headline: the one-walker-per-thread kernel under the 68 B convention ----
    extra = {}
    if spec["det"] and coalesced and headline:
        engine.destroy_index()
        engine.set_config(generate_mode=wharf.GENERATE_PER_WALKER)
        pw = timed_passes(ctx, engine, 2, 1, False)
        pw_s = ctx.max_over_ranks(sum(pw["kernel_ms"]) * 1e-3)
        pw_gbs = (pw["local_steps"] / 2) * BYTES_PER_DET_STEP / (sum(pw["kernel_ms"]) / 2 * 1e-3) / 1e9
        parity["det_fingerprint_per_walker_kernel"] = ctx.wrapping_sum_over_ranks(engine.corpus_fingerprint())
        parity["schedules_equal"] = parity["det_fingerprint_per_walker_kernel"] == parity[fp_key]
        extra["per_walker_kernel"] = {"value": ctx.sum_over_ranks(pw["local_steps"]) / pw_s, "unit": "walk-steps/s", "ms_per_step": 1e3 * pw_s / 2,
                                      "roofline": {"bound": "hbm", "bytes_per_step": BYTES_PER_DET_STEP, "achieved": pw_gbs, "peak": ctx.peak_gbs, "unit": "GB/s",
                                                   "frac": pw_gbs / ctx.peak_gbs, "kernel": "walk_deterministic_ilp_kernel<1,256,0>",
                                                   "note": "SURVEY 8d convention (68 B per walker step) on the kernel it describes: one walker per thread, "
                                                           "two dependent random fetches per step"},
                                      "note": "generate_mode PER_WALKER, 2 timed passes: the same corpus, bit for bit (parity_check.schedules_equal)"}
        engine.destroy_index()
        engine.set_config(generate_mode=wharf.GENERATE_AUTO)
    if not spec["det"]:
        engine.destroy_index()
        engine.set_config(deterministic_mode=True)
        det = timed_passes(ctx, engine, steps, warmup, False)
        det_s = ctx.max_over_ranks(sum(det["kernel_ms"]) * 1e-3)
        det_gbs = (det["local_steps"] / steps) * BYTES_PER_DET_STEP / (sum(det["kernel_ms"]) / steps * 1e-3) / 1e9
        parity["det_fingerprint"] = ctx.wrapping_sum_over_ranks(engine.corpus_fingerprint())
        extra["deterministic_mode"] = {"value": ctx.sum_over_ranks(det["local_steps"]) / det_s, "unit": "walk-steps/s", "ms_per_step": 1e3 * det_s / steps,
                                       "roofline": {"bound": "hbm", "achieved": None, "peak": ctx.peak_gbs, "unit": "GB/s", "frac": None,
                                                    "algorithmic": {"bytes_per_step": BYTES_PER_DET_STEP, "achieved": det_gbs, "over_peak": det_gbs / ctx.peak_gbs},
                                                    "kernel": "walk_deterministic_ilp_kernel<1,256,0>",
                                                    "note": "adjacency L2-resident at this size (desc 8 MB + pool 67 MB): the 68 B convention counts sectors "
                                                            "the L2 serves, so no HBM fraction is claimed (SURVEY 8d caveat); `algorithmic` is the model's figure"},
                                       "note": "config::deterministic_mode=true (reference default): corpus bit-exact vs the reference"}
        engine.destroy_index()
        engine.set_config(deterministic_mode=False)
    golden = ctx.golden_fp.get(name, {})
    for key in ("det_fingerprint", "mh_fingerprint"):
        if key in parity and key in golden:
            parity[key + "_equals_cpu_oracle"] = parity[key] == int(golden[key])
    if golden:
        parity["cpu_oracle_source"] = "tests/golden/bench_fingerprints.json (oracle/wharf_oracle.c on the same generator output; tests/golden/make_bench_fingerprints.py)"

    # ---- e2e through the C ABI with host buffers ----
    e2e = None
    if e2e_iters > 0:
        engine.destroy()
        e2e = run_e2e(ctx, spec, cfg, n, m, d_offsets, d_edges, e2e_iters)
        engine = wharf.WharfMH(n, m, d_offsets.data_ptr(), d_edges.data_ptr(), config=cfg, device_arrays=True)
        if mh_n2v and 1 < world <= 8 and not args.no_peers:
            from dynamicgraphrepresentationlearning_b200.distributed import connect_peers
            connect_peers(engine)

    # ---- edge-batch update latency (second half of BASELINE.json's metric) ----
    upd = None
    if updates:
        search = {"auto": wharf.SEARCH_AUTO, "index": wharf.SEARCH_INDEX, "scan": wharf.SEARCH_SCAN}[args.search]
        engine.set_config(search_mode=search)
        engine.generate_initial_random_walks()
        sizes = [int(x) for x in args.batch_sweep.split(",") if x] or [args.batch_edges] * args.updates
        upd = run_updates(ctx, engine, n, sizes, args.insert_only, args.search)
        if not args.no_index_vs_scan:
            upd["index_vs_scan"] = index_vs_scan(ctx, engine, n, [1000, args.batch_edges])
            upd["auto_rule"] = ("AUTO takes the index when one is valid and the touched vertices' index entries number fewer than 1/8 of the corpus "
                                "cells (engine.cu update()); index_vs_scan times both on the same batch")
    engine.destroy()
    del d_offsets, d_edges
    torch.cuda.empty_cache()
    wharf.trim_memory(local_rank)
    block.update({"roofline": roofline, "e2e": e2e, "updates": upd, "parity_check": parity, "clocks": clocks.summary() if headline else None})
    block.update(extra)
    return block


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="all", choices=["all"] + sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--updates", type=int, default=4, help="edge batches (insert then delete) to time after the walk benchmark; 0 = skip")
    ap.add_argument("--batch-edges", type=int, default=100000)
    ap.add_argument("--batch-sweep", default="", help="comma-separated batch sizes (BASELINE config 5): one insert+delete per size instead of --updates batches")
    ap.add_argument("--insert-only", action="store_true", help="BASELINE config 3: a stream of insert batches, no deletes")
    ap.add_argument("--search", default="auto", choices=["auto", "index", "scan"])
    ap.add_argument("--no-index-vs-scan", action="store_true")
    ap.add_argument("--no-peers", action="store_true", help="N > 1: do not map peer memory (every rank computes all sampler init values itself)")
    args = ap.parse_args()
    if args.warmup < 3:
        log("warmup raised to 3 (timing rules)")
        args.warmup = 3

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import dynamicgraphrepresentationlearning_b200 as wharf

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    ctx = Ctx(args, torch, dist, wharf, np)
    torch.cuda.set_device(ctx.local_rank)
    if ctx.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{ctx.local_rank}"))
    if ctx.world != args.gpus:
        log(f"note: --gpus {args.gpus} but WORLD_SIZE={ctx.world}; using WORLD_SIZE")
    rank, world = ctx.rank, ctx.world

    e2e_iters = 0 if args.no_e2e else max(1, min(args.steps, 3))
    if args.workload == "all":
        head_name = "dw25"
        head = run_workload(ctx, "dw25", steps=args.steps, warmup=args.warmup, e2e_iters=e2e_iters, updates=False, headline=True)
        blocks = {"n2v20": run_workload(ctx, "n2v20", steps=args.steps, warmup=args.warmup, e2e_iters=e2e_iters, updates=args.updates > 0, headline=False),
                  "dw14": run_workload(ctx, "dw14", steps=args.steps, warmup=args.warmup, e2e_iters=e2e_iters, updates=False, headline=False)}
    else:
        head_name = args.workload
        mh = not WORKLOADS[head_name]["det"]
        head = run_workload(ctx, head_name, steps=args.steps, warmup=args.warmup, e2e_iters=e2e_iters, updates=mh and args.updates > 0, headline=True)
        blocks = {}

    # ---- CPU baseline on the box's host cores (rank 0, N = 1 only): the reference itself on BASELINE configs[0], in full ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            _, rspec = reference_spec(args)
            times, steps, cores, kind, sample = cpu_reference_run(rspec, 2, None)
            cpu = {"value": steps / min(times), "unit": "walk-steps/s", "cores": cores, "kind": kind, "sample": sample,
                   "like_for_like": "blocks.dw14.value_with_index (same graph, same 10x80 corpus, inverted index built inside the call on both sides)"
                                    if "dw14" in blocks or head_name == "dw14" else None}
        except Exception as ex:  # noqa: BLE001
            cpu = {"value": None, "unit": "walk-steps/s", "cores": os.cpu_count(), "kind": "unavailable", "sample": repr(ex)}

    if rank == 0:
        spec = WORKLOADS[head_name]
        parity = {head_name: head.pop("parity_check")}
        for k, b in blocks.items():
            parity[k] = b.pop("parity_check")
        parity["note"] = ("wrapping sums over the ranks of wmh_corpus_fingerprint (a sum over walks of the FNV-1a of each walk's text): keyed draws use "
                          "global walk ids, so every N must print the same values")
        line = {
            "metric": "walk-steps/sec (init walks)", "value": head["value"], "unit": "walk-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": spec["desc"], "generator": "utility::generate_batch_of_edges(E/2, V, seed 42, a=.5,b=.2,c=.1, symmetric) restated on device",
                       "n": head["n"], "m": head["m"], "walks_per_vertex": WPV, "walk_length": LEN, "deterministic_mode": spec["det"],
                       "parallelism": f"walks sharded over {world} rank(s) (whole replicas first, the rest by start vertex), graph replicated",
                       "peer_exchanges": head["peer_exchanges"], "l2": head["l2"],
                       "blocks": {k: b["workload"] for k, b in blocks.items()}},
            "clocks": head["clocks"], "gpu_launches": head["gpu_launches"], "wall_s_timed_region": head["wall_s_timed_region"],
            "value_with_index": head.get("value_with_index"), "index_build_ms": head.get("index_build_ms"), "index_note": head.get("index_note"),
            "roofline": head["roofline"], "cpu_baseline": cpu, "e2e": head["e2e"], "updates": head["updates"],
            "parity_check": parity,
        }
        for k in ("per_walker_kernel", "deterministic_mode"):
            if isinstance(head.get(k), dict):
                line[k] = head[k]
        if blocks:
            line["blocks"] = blocks
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
