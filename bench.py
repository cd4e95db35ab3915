#!/usr/bin/env python
"""Headline benchmark of the WharfMH hot path on B200 (contract: see the round brief, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A *step* is one full pass of initial walk generation (graph/wharfmh.h:250) over the workload's graph:
10 walks per vertex of length 80.  `value` = walk-steps/s with the graph resident in HBM, timed with CUDA
events on the engine's stream; `e2e` = the same through the C ABI with host buffers (graph H2D, generate,
corpus D2H inside the timed region).  At N > 1 the walks shard by start vertex over the ranks (graph
replicated, no data-path collective) and the same total work is timed: strong scaling.

Workloads (BASELINE.json configs):
    n2v20  configs[1]: node2vec MH (p=0.5, q=2) on RMAT scale-20 (1M V, 16M E)      <- default
    dw14   configs[0]: DeepWalk on RMAT scale-14 (the reference's CPU-runnable case)
    dw25   configs[3]: DeepWalk on RMAT scale-25 (33M V, 537M E), sharded by start vertex
    lj     configs[2]: LiveJournal-shape RMAT (4.8M V, 69M E) node2vec MH (+ 10 insert batches of 100K)
    n2v23  configs[4]: RMAT scale-23 node2vec MH (batch-size sweep with --updates)
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
    "n2v20": dict(n=1 << 20, vertices=(1 << 20) + 1, edges_half=8 << 20, model="node2vec", det=False, desc="node2vec MH (p=0.5,q=2) 10x80, RMAT scale-20 (1M V, 16M E)"),
    "dw14": dict(n=1 << 14, vertices=(1 << 14) + 1, edges_half=128 << 10, model="deepwalk", det=True, desc="DeepWalk 10x80, RMAT scale-14 (16K V, 256K E)"),
    "dw25": dict(n=1 << 25, vertices=(1 << 25) + 1, edges_half=268435456, model="deepwalk", det=True, desc="DeepWalk 10x80, RMAT scale-25 (33M V, 537M E)"),
    "lj": dict(n=4847571, vertices=4847571, edges_half=34500000, model="node2vec", det=False, desc="node2vec MH 10x80, LiveJournal-shape RMAT (4.8M V, 69M E)"),
    "n2v23": dict(n=1 << 23, vertices=(1 << 23) + 1, edges_half=67108864, model="node2vec", det=False, desc="node2vec MH (p=0.5,q=2) 10x80, RMAT scale-23 (8.4M V, 134M E)"),
}
WPV, LEN, P, Q = 10, 80, 0.5, 2.0
GRAPH_SEED = 42
BYTES_PER_DET_STEP = 68.0      # 32 B row descriptor sector + 32 B neighbour sector + 4 B coalesced store


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ----------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi during the timed region
# ----------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled every 20 ms from a thread of RANK 0 only (the
    timed region of the default run is ~150 ms, one nvidia-smi process launch would not fit into it); nvidia-smi when
    NVML is unavailable.  Other ranks do not poll: NVML queries take driver-wide locks, and eight processes polling
    every few milliseconds showed up as ~1 ms of launch gaps inside a 4 ms multi-launch pass at N = 8."""
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
def cpu_reference_run(spec, trials: int, sample_scale: int):
    """Runs the reference's generate_initial_random_walks `trials` times on a bounded sample of the workload
    (same generator and parameters, smaller scale).  Returns (per-trial seconds, steps per trial, description)."""
    from oracle import wharf_oracle as wo
    n = 1 << sample_scale
    off, adj = wo.rmat_graph(sample_scale, 8 * n, GRAPH_SEED)
    cores = os.cpu_count() or 1
    sample = f"RMAT scale-{sample_scale} ({n} V, {adj.shape[0]} E, same generator/parameters), full 10x80 corpus, {trials} trial(s)"
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


def run_reference_arm(args, spec):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    trials = args.warmup + args.steps
    times, steps, cores, kind, sample = cpu_reference_run(spec, trials, args.cpu_sample_scale)
    timed = times[args.warmup:]
    value = steps * len(timed) / sum(timed)
    line = {
        "impl": "reference", "metric": "walk-steps/sec (init walks)", "value": value, "unit": "walk-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(timed) / len(timed), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": spec["desc"], "walks_per_vertex": WPV, "walk_length": LEN, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "walk-steps/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "walk-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="n2v20", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample-scale", type=int, default=15)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--updates", type=int, default=4, help="edge batches (insert then delete) to time after the walk benchmark; 0 = skip")
    ap.add_argument("--batch-edges", type=int, default=100000)
    ap.add_argument("--batch-sweep", default="", help="comma-separated batch sizes (BASELINE config 5): one insert+delete per size instead of --updates batches")
    ap.add_argument("--insert-only", action="store_true", help="BASELINE config 3: a stream of insert batches, no deletes")
    ap.add_argument("--search", default="auto", choices=["auto", "index", "scan"])
    ap.add_argument("--no-peers", action="store_true", help="N > 1: do not map peer memory (every rank computes all sampler init values itself)")
    args = ap.parse_args()
    spec = WORKLOADS[args.workload]
    if args.warmup < 3:
        log("warmup raised to 3 (timing rules)")
        args.warmup = 3

    if args.impl == "reference":
        run_reference_arm(args, spec)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import dynamicgraphrepresentationlearning_b200 as wharf

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    if world != args.gpus:
        log(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
    dev = torch.device(f"cuda:{local_rank}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: int) -> int:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())

    # ---- input: graph replicated on every rank (generated on the device, identical bytes everywhere) ----
    t0 = time.perf_counter()
    n, m, d_offsets, d_edges = make_graph_on_device(wharf, torch, spec, local_rank)
    torch.cuda.synchronize()
    log(f"[rank {rank}] graph: n={n} m={m} generated in {time.perf_counter() - t0:.2f}s")
    model = wharf.NODE2VEC if spec["model"] == "node2vec" else wharf.DEEPWALK
    search = {"auto": wharf.SEARCH_AUTO, "index": wharf.SEARCH_INDEX, "scan": wharf.SEARCH_SCAN}[args.search]
    cfg = wharf.Config(walks_per_vertex=WPV, walk_length=LEN, random_walk_model=model, paramP=P, paramQ=Q,
                       deterministic_mode=spec["det"], seed=1, search_mode=wharf.SEARCH_SCAN, device=local_rank, rank=rank, world=world)
    engine = wharf.WharfMH(n, m, d_offsets.data_ptr(), d_edges.data_ptr(), config=cfg, device_arrays=True)
    peers = False
    if 1 < world <= 8 and not args.no_peers:
        from dynamicgraphrepresentationlearning_b200.distributed import connect_peers
        peers = connect_peers(engine)                                    # exchange buffers mapped between the ranks (NVLink)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def one_step():
        flush.fill_(1)                                                   # L2 flush between timed iterations
        torch.cuda.synchronize()
        if peers:
            barrier()                                                    # ranks enter the pass together: the peer exchange meets at a flag
                                                                         # barrier ON the device, so host-side skew would be timed as waiting
        steps = engine.generate_initial_random_walks()                   # device-timed inside (CUDA events on the engine stream)
        return steps, engine.last_walk_kernel_ms()

    for _ in range(args.warmup):
        one_step()
    launches0 = engine.kernel_launches()
    barrier()
    with ClockSampler(local_rank, enabled=rank == 0) as clocks:
        wall0 = time.perf_counter()
        kernel_ms, local_steps = [], 0
        for _ in range(args.steps):
            s, ms = one_step()
            local_steps += s
            kernel_ms.append(ms)
        barrier()
        wall = time.perf_counter() - wall0
    launches = engine.kernel_launches() - launches0
    peer_exchanges = engine.peer_exchanges()
    device_s = max_over_ranks(sum(kernel_ms) * 1e-3)
    total_steps = sum_over_ranks(local_steps)
    value = total_steps / device_s
    kernel_avg_ms = sum(kernel_ms) / len(kernel_ms)

    # ---- the inverted index over the corpus (walks/inverted_index.h): derived on demand by one device radix sort; the
    # timed step above stores the corpus as the walk matrix and finds affected walks by scanning it (search_mode SCAN) ----
    index_build_ms = None
    if (n * WPV // world) * LEN < (1 << 32):
        engine.build_index()                                             # first build grows the device memory pool
        torch.cuda.synchronize()
        t_idx = time.perf_counter()
        engine.build_index()
        index_build_ms = max_over_ranks(1e3 * (time.perf_counter() - t_idx))

    # ---- roofline of the dominant kernel (walk kernel), rank 0's launch ----
    peaks = {}
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        peaks = json.loads(peaks_path.read_text())
    peak_gbs, peak_kind = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    if spec["det"]:
        bytes_per_step = BYTES_PER_DET_STEP
    else:
        sample = torch.from_numpy(engine.walks(0, min(20000, engine.number_of_local_walks())).astype(np.int64)).to(dev).t().contiguous()
        bytes_per_step = mh_bytes_per_step(torch, d_offsets, sample) if spec["model"] == "node2vec" else BYTES_PER_DET_STEP
    if spec["det"]:
        coalesced = n * 8 + m * 4 > (96 << 20)          # engine's GENERATE_AUTO rule (engine.cu use_coalesced)
        kernel_name = ("coalesced_{advance,tile_counts,expand,relabel,emit}_kernel segments (kernels_coalesced.cu), one generate pass replayed as a CUDA graph" if coalesced
                       else "walk_deterministic_ilp_kernel<1,256,0>")
    else:
        kernel_name = ("walk_n2v_kernel<false> (+ n2v_share_init_kernel: bulk sampler init, 16 % of the pass)" if spec["model"] == "node2vec"
                       else "walk_deepwalk_mh_kernel<false>")
    steps_per_launch = local_steps / args.steps
    achieved_gbs = steps_per_launch * bytes_per_step / (kernel_avg_ms * 1e-3) / 1e9
    traffic = None
    tpath = ROOT / "profiles" / "roofline_traffic.json"
    if tpath.exists():
        traffic = json.loads(tpath.read_text()).get(f"{args.workload}:{world}")
    roofline = {"bound": "hbm", "achieved": achieved_gbs, "peak": peak_gbs, "unit": "GB/s", "frac": achieved_gbs / peak_gbs,
                "traffic": traffic, "kernel": kernel_name,
                "bytes_per_step": bytes_per_step, "steps_per_launch": steps_per_launch, "kernel_ms": kernel_avg_ms, "peak_source": peak_kind,
                # the bound the north star names: random 32-B sector gathers from HBM, measured with tools/gather_probe.cu
                # (profiles/r01c_gather_probe.txt): 37.7 G sectors/s from a 4 GiB table, whatever the policy or occupancy
                "random_gather": {"peak_gsectors_s": 37.7, "source": "tools/gather_probe.cu, profiles/r01c_gather_probe.txt",
                                  "dram_fetches_per_s_64B": (traffic["read"] / 64 / (traffic.get("kernel_ms", kernel_avg_ms) * 1e-3)) if isinstance(traffic, dict) else None,
                                  "note": "fetch rate = ncu DRAM bytes read / 64 B over the ncu duration of the walk kernel alone"}}
    if isinstance(traffic, dict):
        roofline["traffic"] = traffic["read"] + traffic["write"]
    if spec["det"] and coalesced:
        roofline["note"] = ("achieved = 68 B (SURVEY 8d: one descriptor sector + one neighbour sector + the stored vertex) x walker steps / time; "
                            "frac > 1 because the coalesced schedule fetches a row once for every walker of a replica standing on it "
                            "(walkers of one replica that meet are identical in deterministic mode): traffic = measured DRAM bytes")
    if not spec["det"] and spec["model"] == "node2vec":
        roofline["note"] = ("achieved = SURVEY 8d's algorithmic bytes of a node2vec MH step (descriptors, candidate, sampler state, two binary "
                            "searches of N(prev)) x steps / time; frac > 1 means the kernel moves fewer bytes than that algorithm: the per-edge "
                            "filter answers most membership tests with one sector and half the sectors hit the L2 (traffic = measured DRAM "
                            "bytes). The binding limit is the DRAM fetch rate: see random_gather")

    # ---- deterministic-mode figure beside an MH headline (reference default mode, bit-exact corpus) ----
    extra = {}
    if not spec["det"]:
        engine.destroy_index()
        engine.set_config(deterministic_mode=True)
        for _ in range(args.warmup):
            one_step()
        barrier()
        det_ms, det_steps = [], 0
        for _ in range(args.steps):
            s, ms = one_step()
            det_steps += s
            det_ms.append(ms)
        barrier()
        det_s = max_over_ranks(sum(det_ms) * 1e-3)
        det_total = sum_over_ranks(det_steps)
        det_gbs = (det_steps / args.steps) * BYTES_PER_DET_STEP / (sum(det_ms) / len(det_ms) * 1e-3) / 1e9
        extra["deterministic_mode"] = {"value": det_total / det_s, "unit": "walk-steps/s", "ms_per_step": 1e3 * det_s / args.steps,
                                       "roofline": {"bound": "hbm", "achieved": det_gbs, "peak": peak_gbs, "unit": "GB/s", "frac": det_gbs / peak_gbs,
                                                    "bytes_per_step": BYTES_PER_DET_STEP, "kernel": "walk_deterministic_ilp_kernel<1,256,0>"},
                                       "note": "config::deterministic_mode=true (reference default): corpus bit-exact vs the reference"}
        engine.destroy_index()
        engine.set_config(deterministic_mode=False)

    # ---- e2e through the C ABI with host buffers: graph H2D + generate + corpus D2H inside the timed region ----
    e2e = None
    if not args.no_e2e:
        h_off = torch.empty(n, dtype=torch.int64).pin_memory()
        h_adj = torch.empty(max(m, 1), dtype=torch.int32).pin_memory()
        h_off.copy_(d_offsets[:n])
        h_adj.copy_(d_edges[: max(m, 1)])
        torch.cuda.synchronize()
        engine.destroy()
        w_local = wharf.WharfMH(n, m, h_off.numpy().view(np.uint64), h_adj.numpy().view(np.uint32), config=cfg)
        w_count = w_local.number_of_local_walks()
        w_local.destroy()
        h_walks = torch.empty((w_count, LEN), dtype=torch.int32).pin_memory()
        h_walks_np = h_walks.numpy().view(np.uint32)
        lib = wharf._capi.load()

        e2e_phases = []

        def e2e_step():
            t = [time.perf_counter()]
            eng = wharf.WharfMH(n, m, h_off.numpy().view(np.uint64), h_adj.numpy().view(np.uint32), config=cfg)   # H2D
            t.append(time.perf_counter())
            s = eng.generate_initial_random_walks_to_host(h_walks_np)                                             # generate + D2H, overlapped
            t.append(time.perf_counter())
            eng.destroy()
            t.append(time.perf_counter())
            e2e_phases.append([round(1e3 * (b - a), 2) for a, b in zip(t, t[1:])])
            return s

        e2e_step()
        barrier()
        t_e2e0 = time.perf_counter()
        e_steps = 0
        e2e_iters = max(1, min(args.steps, 3))
        for _ in range(e2e_iters):
            e_steps += e2e_step()
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t_e2e0)
        e2e = {"value": sum_over_ranks(e_steps) / e2e_s, "unit": "walk-steps/s",
               "h2d_bytes_per_step": int(n * 8 + m * 4), "d2h_bytes_per_step": int(w_count * LEN * 4),
               "iters": e2e_iters, "ms_per_step": 1e3 * e2e_s / e2e_iters,
               "phases_ms_last": dict(zip(("create_h2d", "generate_to_host", "destroy"), e2e_phases[-1])),
               "path": "wmh_create(host CSR) + wmh_generate_initial_random_walks_to_host(pinned host: whole corpus D2H overlapped with the walk kernels) + wmh_destroy, per rank"}
        engine = wharf.WharfMH(n, m, d_offsets.data_ptr(), d_edges.data_ptr(), config=cfg, device_arrays=True)

    # ---- edge-batch update latency (second half of BASELINE.json's metric) ----
    updates = None
    if args.updates > 0:
        engine.set_config(search_mode=search)
        engine.generate_initial_random_walks()
        if search == wharf.SEARCH_AUTO:
            pass
        rows = []
        sizes = [int(x) for x in args.batch_sweep.split(",") if x] or [args.batch_edges] * args.updates
        for b, size in enumerate(sizes):
            d_batch, k = make_batch_on_device(wharf, torch, size, n, b, local_rank)
            h_batch = torch.empty(2 * k, dtype=torch.int32).pin_memory()
            h_batch.copy_(d_batch)
            torch.cuda.synchronize()
            hb = h_batch.numpy().view(np.uint32).reshape(-1, 2)
            row = {"batch_edges": k}
            for kind in (("insert",) if args.insert_only else ("insert", "delete")):
                engine.reset_timers()
                barrier()
                t0 = time.perf_counter()
                fn = engine._lib.wmh_insert_edges_batch if kind == "insert" else engine._lib.wmh_delete_edges_batch
                na = C.c_uint64()
                wharf._capi.check(fn(engine._h, k, C.c_void_p(hb.ctypes.data), 1, 0, (1 << 64) - 1, 1, C.byref(na)))
                barrier()
                dt = max_over_ranks(time.perf_counter() - t0)
                gi = 0 if kind == "insert" else 2
                row[kind] = {"latency_ms": 1e3 * dt, "graph_update_ms": 1e3 * max_over_ranks(engine.timer(gi)),
                             "walk_update_ms": 1e3 * max_over_ranks(engine.timer(gi + 1)), "affected_walks": sum_over_ranks(int(na.value)),
                             "rewalk_steps": sum_over_ranks(engine.last_steps()), "rewalk_kernel_ms": max_over_ranks(engine.last_walk_kernel_ms())}
            rows.append(row)
        ins = [r["insert"]["latency_ms"] for r in rows]
        dele = [r["delete"]["latency_ms"] for r in rows if "delete" in r] or [float("nan")]
        updates = {"search": args.search, "batches": rows, "insert_latency_ms_median": statistics.median(ins),
                   "delete_latency_ms_median": statistics.median(dele), "host_batch": True,
                   "note": "latency = host call of wmh_insert/delete_edges_batch (H2D batch, merge, affected search, re-walk), max over ranks"}

    # ---- CPU baseline on the box's host cores (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            times, steps, cores, kind, sample = cpu_reference_run(spec, 1, args.cpu_sample_scale)
            cpu = {"value": steps / times[0], "unit": "walk-steps/s", "cores": cores, "kind": kind, "sample": sample}
        except Exception as ex:  # noqa: BLE001
            cpu = {"value": None, "unit": "walk-steps/s", "cores": os.cpu_count(), "kind": "unavailable", "sample": repr(ex)}

    if rank == 0:
        line = {
            "metric": "walk-steps/sec (init walks)", "value": value, "unit": "walk-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * device_s / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": spec["desc"], "generator": "utility::generate_batch_of_edges(E/2, V, seed 42, a=.5,b=.2,c=.1, symmetric) restated on device",
                       "n": n, "m": m, "walks_per_vertex": WPV, "walk_length": LEN, "deterministic_mode": spec["det"],
                       "parallelism": f"walks sharded by start vertex over {world} rank(s), graph replicated"
                                      + (", sampler init values computed once per box and stored into the peers over NVLink" if peers else ""),
                       "peer_exchanges": peer_exchanges,
                       "l2": "explicit 256 MiB write between timed steps"},
            "clocks": clocks.summary(), "gpu_launches": launches, "wall_s_timed_region": wall,
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "updates": updates,
            "index_build_ms": index_build_ms,
        }
        line.update(extra)
        print(json.dumps(line), flush=True)
    engine.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
