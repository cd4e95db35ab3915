"""Turns gpurun_out/launches*.csv (ncu --metrics gpu__time_duration.sum) and gpurun_out/*.ncu-rep (ncu --set full)
into the text summaries committed under profiles/.  Usage:
    python profiles/summarise.py launches gpurun_out/launches.csv > profiles/rNN_launches.txt
    python profiles/summarise.py kernel   gpurun_out/prof_walk.ncu-rep > profiles/rNN_walk_kernel.txt
    python profiles/summarise.py traffic  gpurun_out/traffic.csv [key] > profiles/rNN_traffic.txt
        (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum: DRAM bytes per kernel of one pass;
         with a key such as dw25:8 the totals are also written into profiles/roofline_traffic.json)
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_ld.ratio",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "sm__maximum_warps_per_active_cycle_pct", "smsp__inst_executed.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = [r for r in csv.reader(lines) if len(r) > 5]
    hdr = rows[0]
    i_name, i_val, i_unit = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[i_val].replace(",", ""))
        except ValueError:
            continue
        ms = {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(r[i_unit], 1e-6) * v
        agg[r[i_name][:110]][0] += 1
        agg[r[i_name][:110]][1] += ms
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot:.3f} ms of kernel time (serialised, cold cache: compare shares)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]:11.3f} ms {100 * v[1] / tot:6.2f}%  x{v[0]:<4d} {k}")


def kernel(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        if "nan" in r[idx["dram__bytes_read.sum"]]:
            continue
        print(f"## {r[idx['Kernel Name']][:140]}")
        for k in KEYS:
            if k in idx:
                print(f"{k:90s} {r[idx[k]]:>22s} {units[idx[k]]}")
        print()


def traffic(path, key=None):
    import json
    import pathlib
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = [r for r in csv.reader(lines) if len(r) > 5]
    hdr = rows[0]
    i_id, i_name, i_metric, i_val, i_unit = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    agg = collections.defaultdict(lambda: {"n": set(), "ms": 0.0, "read": 0.0, "write": 0.0})
    for r in rows[1:]:
        try:
            v = float(r[i_val].replace(",", "")) * scale.get(r[i_unit], 1.0)
        except ValueError:
            continue
        a = agg[r[i_name][:100]]
        a["n"].add(r[i_id])
        if r[i_metric] == "gpu__time_duration.sum":
            a["ms"] += v
        elif r[i_metric] == "dram__bytes_read.sum":
            a["read"] += v
        elif r[i_metric] == "dram__bytes_write.sum":
            a["write"] += v
    ms = sum(a["ms"] for a in agg.values())
    rd = sum(a["read"] for a in agg.values())
    wr = sum(a["write"] for a in agg.values())
    print(f"# {path}: {sum(len(a['n']) for a in agg.values())} launches; per-launch times are serialised and cold-cache (compare shares)")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        gbs = (a["read"] + a["write"]) / a["ms"] / 1e6 if a["ms"] else 0.0
        print(f"{a['ms']:9.3f} ms x{len(a['n']):<4d} read {a['read'] / 1e9:8.2f} GB  write {a['write'] / 1e9:8.2f} GB  {gbs:7.0f} GB/s  {k}")
    print(f"# total: {ms:.3f} ms serialised, read {rd / 1e9:.2f} GB + write {wr / 1e9:.2f} GB = {(rd + wr) / 1e9:.2f} GB per pass")
    if key:
        f = pathlib.Path(__file__).with_name("roofline_traffic.json")
        data = json.loads(f.read_text())
        data[key] = {"read": rd, "write": wr, "pass_kernel_ms_serialised": round(ms, 3),
                     "source": f"ncu dram__bytes_read.sum + dram__bytes_write.sum over the kernels of one generate pass of rank 0's shard (tools/profile_pass.py "
                               f"{key.split(':')[0]} <mode> 0 {key.split(':')[1]} on one GPU; MH shards without the peer exchange: every sampler init is "
                               f"computed locally, an upper bound of a rank's traffic)"}
        f.write_text(json.dumps(data, indent=1) + "\n")


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
