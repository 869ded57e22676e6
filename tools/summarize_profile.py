"""Turn gpurun_out ncu artefacts into the tracked summaries under profiles/.

usage: python tools/summarize_profile.py <tag> <launches.csv> <full.ncu-rep> [workload] [n_gpus]
       python tools/summarize_profile.py --traffic <workload> <n_gpus> <dram.csv>     (a --metrics dram__bytes_* capture of one sweep launch)
writes profiles/<tag>_launches.md, profiles/<tag>_sweep_ncu.md and the entry [workload][n_gpus] of profiles/sweep_traffic.json
"""
import collections, csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__registers_per_thread",
        "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]


WORKLOAD, NGPUS = "c4", "1"


def set_traffic(workload, n, value, source):
    """profiles/sweep_traffic.json: {workload: {n_gpus: dram bytes of ONE sweep launch of one rank}, "_source": {...}}"""
    path = os.path.join(ROOT, "profiles", "sweep_traffic.json")
    try:
        d = json.load(open(path))
        if "dram_bytes_per_launch" in d:
            d = {}
    except Exception:
        d = {}
    d.setdefault(workload, {})[str(n)] = value
    d.setdefault("_source", {})[f"{workload}/{n}"] = source
    json.dump(d, open(path, "w"), indent=1, sort_keys=True)


def traffic_csv(workload, n, path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, ui, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for r in rows[1:]:
        if "sweep_kernel" in r[ki] and r[mi].startswith("dram__bytes_"):
            tot += float(r[vi].replace(",", "")) * mult[r[ui]]
    set_traffic(workload, n, tot, os.path.basename(path))
    print(workload, n, tot)


def launches(tag, path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        k = r[ki].split("(")[0][:70]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values())
    out = [f"# {tag}: ncu launch list (`--metrics gpu__time_duration.sum --clock-control none`)", "",
           "Per-launch times are cold-cache and serialised under ncu: compare SHARES, not absolutes.", "",
           "| kernel | launches | total us | us/launch | share |", "|---|---|---|---|---|"]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| `{k}` | {a[0]} | {a[1] / 1e3:.1f} | {a[1] / 1e3 / a[0]:.1f} | {a[1] / tot:.3f} |")
    open(os.path.join(ROOT, "profiles", f"{tag}_launches.md"), "w").write("\n".join(out) + "\n")


def full(tag, rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = [f"# {tag}: `ncu --set full --clock-control none` of the dominant kernel", ""]
    for vals in rows[2:]:
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        out += [f"## {name[:100]}", "", "| metric | unit | value |", "|---|---|---|"]
        d = dict(zip(hdr, vals))
        u = dict(zip(hdr, units))
        for k in KEYS:
            if k in d:
                out.append(f"| {k} | {u[k]} | {d[k]} |")
        try:
            mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tr = float(d["dram__bytes_read.sum"]) * mult[u["dram__bytes_read.sum"]] + \
                float(d["dram__bytes_write.sum"]) * mult[u["dram__bytes_write.sum"]]
            if "sweep_kernel" in name:
                set_traffic(WORKLOAD, NGPUS, tr, f"profiles/{tag}_sweep_ncu.md")
            out.append(f"| dram read+write per launch | byte | {tr:.0f} |")
        except Exception:
            pass
        out.append("")
    open(os.path.join(ROOT, "profiles", f"{tag}_sweep_ncu.md"), "w").write("\n".join(out) + "\n")


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":
        traffic_csv(sys.argv[2], sys.argv[3], sys.argv[4])
        sys.exit(0)
    tag = sys.argv[1]
    if len(sys.argv) > 5:
        WORKLOAD, NGPUS = sys.argv[4], sys.argv[5]
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    launches(tag, sys.argv[2])
    if len(sys.argv) > 3:
        full(tag, sys.argv[3])
