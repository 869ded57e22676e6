#!/usr/bin/env python
"""bench.py -- (tx,rx) SINR link decisions per second per tick (BASELINE.json metric), SURVEY.md 8(d).

Workloads (`--workload`, inputs from java.util.Random -- tools/workloads.py -- so a JVM regenerates them bit for bit):
  c4 (default)  BASELINE configs[3], the tick the metric and north_star are quoted on: 1 000 000 nodes uniform in a
                10 km square, 10 000 concurrent transmitters, one channel, log-distance path loss + co-channel
                interference + SINR decisions.  It fits one B200, so the same tick is used at every N; at N > 1 the
                receivers are block-partitioned over the ranks (strong scaling, as north_star asks).
  c4-grid       the same on the 1000 x 1000 grid (10 m pitch)
  c2, c3        BASELINE configs[1], [2] (1k x 50, 100k x 1k): latency lines -- a tick of 5e4 pairs is launch-bound
  c1            BASELINE configs[0]: the 10-node grid over the JSON radio protocol (UDGM), one broadcast per tick
  c5            BASELINE configs[4]: 1 M nodes on 16 channels, every node moves every tick, fresh transmitters per tick
                (`--steps 1000` = the configuration's 1000 ticks)
Variants of c2..c4: `--pl-exponent 2.7` (general exponent: lg2 + ex2 = 2 MUFU per decision), `--tx-power -25|-15|-7|0`
(CC2420 levels), `--rx-prob 0.9` (stochastic reception, uniforms supplied by the host from java.util.Random; N = 1).

A "step" is one tick; steps cycle through up to 8 different transmitter lists.  JSON line:
  value     decisions / device time of the tick, tx records resident in HBM, results left in HBM.  CUDA events on the
            stream the library runs on; L2 flush (256 MiB write), ranks lined up ON THE DEVICE (rmb_peer_barrier_enqueue),
            start event, tick, stop event enqueued back to back; per-step times summed, max over ranks.
  e2e       the same tick through the public C-ABI call with HOST packet lists in and HOST result lists out (compact
            form: what the reference-side adapter consumes), host wall clock.  THE HEADLINE against --impl reference.
  roofline  dominant kernel = sweep_kernel, bound = the XU (MUFU) pipe; `achieved` = MUFU ops per launch / the launch's
            duration by CUDA events in a pass of its own (events BETWEEN the tick's kernels serialise them, so the
            value/e2e legs run without; they report the kernels' own time stamps as `ms_per_launch_device_stamps`);
            `peak` = 148 SM x 16 lanes x sm_max_mhz, beside it what radio-sim_b200/mufu_peak measures on this GPU.
  cpu_baseline  the oracle port (plain C restatement of the medium's loop) on the box's host cores: all threads AND
            one thread (the reference's loop is single-threaded, UDGMRadioMedium.java:99), warmed, median of 5, on a
            bounded receiver sample of the same tick.
  verified  the tick's (tx, rx, outcome) lists digested and compared with tests/golden/fullsize_hashes.json (the
            oracle's lists); at N > 1 after the per-rank lists were merged.

`--impl reference`: the reference is Java and cannot be built here (no JVM), so this arm times the oracle port on
all host threads, each step a bounded sample of the same workload; rank 0 only.
`--single-process`: N handles in ONE process, one host thread per GPU, the per-rank result lists all in this
process's host memory inside the timed region (what one JVM owning N handles sees).
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import numpy as np

METRIC = "tx_rx_sinr_link_decisions_per_sec"
UNIT = "decisions/s"
SMS, XU_LANES = 148, 16
WORKLOAD_TEXT = {
    "c1": "10-node 802.15.4 grid, unit-disk medium, one broadcast per node over the JSON radio protocol (BASELINE configs[0])",
    "c2": "1k-node random deployment, log-distance path loss, 50 concurrent transmitters, SINR collisions (BASELINE configs[1])",
    "c3": "100k-node dense deployment, 1k concurrent transmitters, single channel, full interference (BASELINE configs[2])",
    "c4": "1M-node random deployment (10 km square), 10k concurrent transmitters, single channel, SINR (BASELINE configs[3])",
    "c4-grid": "1M-node grid deployment (1000 x 1000, 10 m pitch), 10k concurrent transmitters, single channel, SINR (BASELINE configs[3], grid)",
    "c5": "1M-node 16-channel 802.15.4 deployment, per-tick mobility of every node, 10k fresh transmitters per tick (BASELINE configs[4])",
}
CPU_SAMPLE_ALL, CPU_SAMPLE_ONE = 50_000, 5_000          # receivers in the cpu_baseline samples (all threads / one thread)


# ---------------------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period=0.005):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.active, self.period = False, period

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop_flag:
                if self.active:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for k, bit in names.items():
                        if r & bit:
                            self.reasons.add(k)
                time.sleep(self.period)
        except Exception as e:          # NVML missing: report that rather than invent clocks
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def digest(tx, rx, outcome):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(tx, dtype="<i4").tobytes())
    h.update(np.ascontiguousarray(rx, dtype="<i4").tobytes())
    h.update(np.ascontiguousarray(outcome, dtype="u1").tobytes())
    return h.hexdigest()


def golden(key):
    try:
        return json.load(open(os.path.join(ROOT, "tests", "golden", "fullsize_hashes.json"))).get(key)
    except Exception:
        return None


def mufu_microbench():
    """radio-sim_b200/mufu_peak (csrc/mufu_peak.cu): what the XU and FMA pipes of this GPU sustain."""
    exe = os.path.join(ROOT, "radio-sim_b200", "mufu_peak")
    try:
        return json.loads(subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout)
    except Exception as e:
        return {"error": f"{type(e).__name__}"}


def mufu_per_decision(n):
    return 1 if n in (2.0, 3.0, 4.0) else 2           # rcp / rsqrt / rcp, else lg2 + ex2 (csrc/sweep.cu)


def traffic_for(workload, gpus):
    """dram__bytes_read + write of ONE sweep launch from an `ncu --set full` capture of this workload at this shard
    size (profiles/sweep_traffic.json, one entry per workload and N); None = never captured."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "sweep_traffic.json")))[workload][str(gpus)]
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------------------------
# inputs
def source_lists(dep, count):
    """list 0 = the configuration's own transmitters; the others are drawn the same way from later seeds"""
    from radio_sim_b200.javarandom import JavaRandom
    out = [dep["src"]]
    for i in range(1, count):
        out.append(JavaRandom(1000 * dep["seed"] + i).sample_distinct(dep["n"], dep["t"]))
    return out


def node_arrays(dep, args):
    n = dep["n"]
    txp = np.full(n, float(args.tx_power)) if args.tx_power is not None else None
    rxp = np.full(n, float(args.rx_prob)) if args.rx_prob is not None else None
    return txp, rxp


def medium_params(args):
    return {"pl_exponent": float(args.pl_exponent)} if args.pl_exponent is not None else {}


# ---------------------------------------------------------------------------------------------------------------
# CPU arm (oracle port)
def cpu_sample(dep, src, args, threads, receivers, reps=5, warm=1):
    """The oracle on a bounded sample of the tick: all packets x the first `receivers` nodes (the transmitters are kept
    as nodes -- they must exist -- but only count as receivers when they fall into the sample)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc
    n = dep["n"]
    receivers = min(receivers, n)
    keep = np.zeros(n, dtype=bool)
    keep[:receivers] = True
    en = keep.copy()
    keep[src] = True
    idx = np.flatnonzero(keep)
    remap = -np.ones(n, dtype=np.int64)
    remap[idx] = np.arange(idx.size)
    txp, rxp = node_arrays(dep, args)
    nodes = orc.NodeSet(dep["x"][idx], dep["y"][idx], channel=dep["channel"][idx], enabled=en[idx].astype(np.uint8),
                        rxprob=None if rxp is None else rxp[idx])
    power = 0.0 if txp is None else float(args.tx_power)
    pk = orc.make_packets(remap[src].astype(np.int32), power, dep["channel"][src])
    p = orc.params(orc.MODEL_SINR, **medium_params(args))
    u = orc.jrandom_doubles(7, 2_000_000) if rxp is not None else None
    times, dec = [], 0
    for i in range(warm + reps):
        t0 = time.perf_counter()
        r = orc.tick(p, nodes, pk, uniforms=u, threads=threads, want_pairs=True, cap=4_000_000)
        dt = time.perf_counter() - t0
        dec = r["n_decisions"]
        if i >= warm:
            times.append(dt)
    return dec, times, f"{receivers} receivers (the first nodes of the deployment) x {len(src)} packets of the same tick"


def cpu_baseline(dep, src, args, quick=False):
    cores = host_threads()
    n_all, n_one = (CPU_SAMPLE_ALL, CPU_SAMPLE_ONE) if dep["n"] * len(src) > 5e8 else (dep["n"], max(1, dep["n"] // 10))
    if dep.get("channels_used", 1) > 1:
        n_all, n_one = n_all * 8, n_one * 8                                  # 1/16 of the packets per receiver
    reps = 2 if quick else 5
    dec_a, t_a, s_a = cpu_sample(dep, src, args, cores, n_all, reps=reps)
    dec_1, t_1, s_1 = cpu_sample(dep, src, args, 1, n_one, reps=reps)
    return {"value": dec_a / float(np.median(t_a)), "unit": UNIT, "cores": cores, "kind": "port", "sample": s_a,
            "stat": f"median of {reps} after 1 warm-up",
            "single_thread": {"value": dec_1 / float(np.median(t_1)), "unit": UNIT, "cores": 1, "sample": s_1,
                              "note": "the reference's transmit() loop is single-threaded (UDGMRadioMedium.java:99)"}}


def run_reference(args):
    """--impl reference: rank 0 times the CPU port of the reference loop; other ranks exit 0."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import workloads as W
    name = args.workload
    if name == "c1":
        return run_c1_reference(args)
    dep = W.deployment("c4" if name == "c4-grid" else name, grid=name == "c4-grid")
    if name == "c5":
        dep["channels_used"] = 16
        src = W.c5_tick(dep, 1)
    else:
        src = dep["src"]
    cores = host_threads()
    recv = CPU_SAMPLE_ALL * (8 if name == "c5" else 1) if dep["n"] * len(src) > 5e8 else dep["n"]
    dec, times, sample = cpu_sample(dep, src, args, cores, recv, reps=args.steps, warm=args.warmup)
    total = sum(times)
    value = dec * len(times) / total
    one = cpu_sample(dep, src, args, 1, max(1, recv // 10), reps=3)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": dict(config_of(args, dep), sample=sample),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "single_thread": {"value": one[0] / float(np.median(one[1])), "unit": UNIT, "cores": 1, "sample": one[2]}},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def config_of(args, dep, **extra):
    c = {"workload": WORKLOAD_TEXT[args.workload], "nodes": int(dep["n"]), "transmitters": int(dep["t"]),
         "inputs": "java.util.Random (tools/workloads.py)"}
    if args.pl_exponent is not None:
        c["pl_exponent"] = float(args.pl_exponent)
    if args.tx_power is not None:
        c["tx_power_dbm"] = float(args.tx_power)
    if args.rx_prob is not None:
        c["rx_probability"] = float(args.rx_prob)
    c.update(extra)
    return c


# ---------------------------------------------------------------------------------------------------------------
def setup_dist():
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the medium has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    return rank, world, local, dev


def open_peers(med, G, rank, max_slots, dev):
    """map every rank's gather buffer (CUDA IPC): the 128-byte blobs travel through one NCCL all-gather at setup"""
    import torch
    import torch.distributed as dist
    blob = torch.frombuffer(bytearray(med.peer_export(G, rank, max_slots)), dtype=torch.uint8).to(dev)
    blobs = torch.empty(G * blob.numel(), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(blobs, blob)
    raw = blobs.cpu().numpy().tobytes()
    med.peer_open([raw[i * blob.numel():(i + 1) * blob.numel()] for i in range(G)])


def run_static(args):
    """c2 / c3 / c4 / c4-grid: static nodes, one channel."""
    import torch
    import torch.distributed as dist
    import radio_sim_b200 as rsb
    from radio_sim_b200.sharding import shard_bounds, TickPlan, merge_results
    from radio_sim_b200.javarandom import JavaRandom
    import workloads as W

    rank, G, local, dev = setup_dist()
    name = args.workload
    dep = W.deployment("c4" if name == "c4-grid" else name, grid=name == "c4-grid")
    N, T = dep["n"], dep["t"]
    stochastic = args.rx_prob is not None and args.rx_prob < 1.0
    if stochastic and G > 1:
        raise SystemExit("--rx-prob < 1 is benchmarked at N = 1 (sharded draws need the rmb_tick_draws_begin/_end pair)")
    bounds = shard_bounds(N, G)
    lo, hi = bounds[rank]
    lists = source_lists(dep, max(1, min(args.steps, 8)))
    plans = [TickPlan(s, bounds) for s in lists]
    tpad_max = max(p.tpad for p in plans)
    txp, rxp = node_arrays(dep, args)
    tuning = args.variant & 0xffffff
    # compact result form: what the reference-side adapter consumes (rx, rssi, outcome per row + first row per packet)
    med = rsb.Medium("sinr", device=local, rx_base=lo, n_shards=G, compact=True, **medium_params(args))
    med.set_nodes(dep["x"][lo:hi], dep["y"][lo:hi], txpower_dbm=None if txp is None else txp[lo:hi],
                  rxprob=None if rxp is None else rxp[lo:hi])
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)                   # NCCL collectives and the library share this stream
    med.set_stream(stream.cuda_stream)
    if tuning:
        med.set_tuning(tuning, 0)
    p2p = G > 1 and args.exchange == "p2p"
    if p2p:
        open_peers(med, G, rank, tpad_max, dev)

    # tx records of every list: this rank's share (padded), resident in HBM before any timed step
    rec_local = [torch.zeros(max(p.tpad, 1) * 48, dtype=torch.uint8, device=dev) for p in plans]
    for p, rl in zip(plans, rec_local):
        med.build_tx_records(p.local_src[rank], rl.data_ptr())
    rec_all = torch.zeros(G * tpad_max * 48, dtype=torch.uint8, device=dev) if G > 1 else None
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    align_token = torch.zeros(1, device=dev)
    uniforms = JavaRandom(7).next_doubles(2_000_000) if stochastic else None

    def line_up():
        if G > 1:
            if p2p:
                med.peer_barrier_enqueue()          # device-side rendezvous: the ranks leave within an NVLink round trip
            else:
                dist.all_reduce(align_token)

    def step_resident(k):
        p = plans[k]
        if stochastic:
            return med.tick_records(p.n_records, rec_local[k].data_ptr(), uniforms=uniforms, copy=False)
        if p2p:
            med.peer_exchange_tick_enqueue(p.tpad, rec_local[k].data_ptr())      # exchange fused into the tick's first kernel
        elif G > 1:
            dist.all_gather_into_tensor(rec_all[:G * p.tpad * 48], rec_local[k][:p.tpad * 48])
            med.tick_enqueue(p.n_records, rec_all.data_ptr())
        else:
            med.tick_enqueue(p.n_records, rec_local[k].data_ptr())
        return None

    def step_e2e(k):
        p = plans[k]
        if G == 1:
            return med.tick(lists[k], uniforms=uniforms, copy=False)
        if p2p:
            return med.peer_tick(p.local_src[rank], copy=False)
        med.build_tx_records(p.local_src[rank], rec_local[k].data_ptr())
        dist.all_gather_into_tensor(rec_all[:G * p.tpad * 48], rec_local[k][:p.tpad * 48])
        return med.tick_records(p.n_records, rec_all.data_ptr(), copy=False)

    def timed(fn, steps, sampler=None, device_timed=True):
        times, walls, sweep_ms, launches, dec, pairs, phases = [], [], [], 0, 0, 0, []
        for i in range(steps):
            k = i % len(lists)
            torch.cuda.synchronize()
            if G > 1:
                dist.barrier()
                torch.cuda.synchronize()
            if sampler:
                sampler.active = True
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush_buf.fill_(1)                      # evict L2; the host runs ahead while the flush executes
            if device_timed:
                line_up()
            else:
                # host-timed leg: the ranks' HOSTS are released together -- a device-side rendezvous, then each host
                # waits for its own GPU (a collective's barrier lets the ranks go tens of microseconds apart, and one
                # rank's late start is every other rank's wait for its records)
                line_up()
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            if device_timed:
                e0.record(stream)
                res = fn(k)
                e1.record(stream)
                torch.cuda.synchronize()
                walls.append(time.perf_counter() - t0)
                if res is None:
                    res = med.tick_finish()
                times.append(e0.elapsed_time(e1) * 1e-3)
            else:
                # the call is synchronous: when it returns the lists are in host memory -- that is the end of the step
                res = fn(k)
                walls.append(time.perf_counter() - t0)
                times.append(walls[-1])
            if sampler:
                sampler.active = False
            tm = med.timing()
            sweep_ms.append(tm["ms_sweep"]); launches += tm["kernel_launches"]
            phases.append((tm["ms_prepare"], tm["ms_sweep"], tm["ms_emit"], tm["ms_total"], times[-1] * 1e3))
            dec += res.n_decisions; pairs += res.n_pairs
        if os.environ.get("RMB_BENCH_DEBUG") and device_timed and steps > 3:
            print(f"[rank {rank}] median ms: prepare(+peer wait) %.4f sweep %.4f rest %.4f total(lib events) %.4f step(events) %.4f" %
                  tuple(np.median(np.array(phases), axis=0)), file=sys.stderr, flush=True)
        return dict(times=times, walls=walls, sweep_ms=sweep_ms, launches=launches, dec=dec, pairs=pairs, last=res)

    sampler = ClockSampler(local, period=0.0003 if N * T < 5e8 else 0.005)
    sampler.start()
    timed(step_resident, args.warmup)
    rv = timed(step_resident, args.steps, sampler)
    # the roofline's kernel duration by CUDA events: a pass of its own, events between the tick's kernels
    med.set_tuning(tuning | (1 << 24), 0)
    timed(step_resident, 2)
    rk = timed(step_resident, args.steps)
    med.set_tuning(tuning, 0)
    timed(step_e2e, args.warmup, device_timed=False)
    re_ = timed(step_e2e, args.steps, device_timed=False)
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # verification: list 0 once more through the host-list path, merged over the ranks, digest vs the oracle's
    verified = None
    gold = golden(name) if not (args.pl_exponent or args.tx_power is not None or args.rx_prob is not None) else None
    if not args.no_verify:
        r0 = step_e2e(0)
        part = dict(tx=r0.tx_column(), rx=r0.rx.copy(), outcome=r0.outcome.copy())
        if G > 1:
            parts = [None] * G
            dist.all_gather_object(parts, part)
        else:
            parts = [part]
        if rank == 0:
            if G > 1:
                m = merge_results(plans[0], parts)
            else:
                m = part
            d = digest(m["tx"], m["rx"], m["outcome"])
            order_ok = bool(np.all((np.diff(m["tx"]) > 0) | ((np.diff(m["tx"]) == 0) & (np.diff(m["rx"]) > 0))))
            verified = {"heard_pairs": int(len(m["rx"])), "sha256": d, "order_ok": order_ok,
                        "golden": None if gold is None else bool(gold["sha256"] == d and gold["n_pairs"] == len(m["rx"])),
                        "what": "tx, rx, outcome lists of transmitter list 0" + (" merged over the ranks" if G > 1 else "")
                                + " vs tests/golden/fullsize_hashes.json (the oracle's lists)"}

    tot = torch.tensor([sum(rv["times"]), sum(re_["walls"]), sum(rv["walls"])], dtype=torch.float64, device=dev)
    cnt = torch.tensor([float(rv["dec"]), float(re_["pairs"]), float(re_["dec"])], dtype=torch.float64, device=dev)
    if G > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    t_res, t_e2e, t_wall = float(tot[0]), float(tot[1]), float(tot[2])
    decisions, pairs_e2e, dec_e2e = float(cnt[0]), float(cnt[1]), float(cnt[2])
    if rank == 0:
        clocks = sampler.summary()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        f_max = (clocks["sm_max_mhz"] or peaks.get("sm_max_mhz") or 1965.0) * 1e6
        f_run = (clocks["sm_mhz"] or 0) * 1e6
        mpd = mufu_per_decision(float(args.pl_exponent) if args.pl_exponent is not None else 3.0)
        ms_sweep_ev = float(np.mean(rk["sweep_ms"]))
        ms_sweep_st = float(np.mean(rv["sweep_ms"]))
        local_dec = rv["dec"] / args.steps                                    # this GPU, per launch
        achieved = mpd * local_dec / (ms_sweep_ev * 1e-3)
        peak = SMS * XU_LANES * f_max
        pairs_tick = pairs_e2e / args.steps
        hbm_bytes = 56.0 * (hi - lo) + 29.0 * pairs_tick / G + 32.0 * T
        launch_bound = N * T < 5e8
        micro = mufu_microbench()
        line = {"metric": METRIC, "value": decisions / t_res, "unit": UNIT, "n_gpus": G,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_res / args.steps,
                "us_per_tick": 1e6 * t_res / args.steps, "ms_per_step_wall": 1e3 * t_wall / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32+f64",
                "data": "synthetic",
                "config": config_of(args, dep, deployment="grid" if name == "c4-grid" else "random",
                                    parallelism=f"receivers sharded x{G}", transmitter_lists=len(lists),
                                    exchange="none" if G == 1 else ("peer-memory stores + flags inside the tick's first kernel" if p2p else "NCCL all-gather"),
                                    l2="flushed (256 MiB write) before every timed step",
                                    timing="value: CUDA events on the library's stream, per step, summed, max over ranks; L2 flush"
                                           + ("" if G == 1 else ", device-side rendezvous of the ranks,")
                                           + " start event, tick and stop event enqueued back to back; e2e: host wall clock around "
                                             "the public C-ABI call with host buffers, max over ranks"),
                "clocks": clocks,
                "e2e": {"value": dec_e2e / t_e2e, "unit": UNIT, "ms_per_step": 1e3 * t_e2e / args.steps,
                        "h2d_bytes_per_step": int(4 * (T if G == 1 else tpad_max) + (8 * 2_000_000 if stochastic else 0)),
                        "d2h_bytes_per_step": int(9 * pairs_tick / G + 5 * (T if G == 1 else G * tpad_max) + 4 + 96),
                        "heard_pairs": int(pairs_tick),
                        "result_form": "compact (RMB_FLAG_COMPACT): int32 rx + float rssi + uint8 outcome per heard pair, int32 first row per packet"},
                "gpu_launches": int(rv["launches"]),
                "roofline": {"bound": "launch latency (a tick of this size is a few dependent launches; no throughput claim)" if launch_bound else "sfu",
                             "kernel": "sweep_kernel", "achieved": achieved / 1e9, "peak": peak / 1e9,
                             "unit": f"G MUFU-op/s ({mpd} per decision; peak = 148 SM x 16 lanes x sm_max_mhz)",
                             "frac": achieved / peak,
                             "frac_at_sampled_clock": (achieved / (SMS * XU_LANES * f_run)) if f_run else None,
                             "ms_per_launch": ms_sweep_ev, "ms_per_launch_device_stamps": ms_sweep_st,
                             "ms_per_step_with_kernel_events": 1e3 * sum(rk["times"]) / args.steps,
                             "traffic": traffic_for(name, G),
                             "measured_pipe_peaks": micro,
                             "hbm": {"achieved_gbs": hbm_bytes / (ms_sweep_ev * 1e-3) / 1e9, "peak_gbs": peaks.get("hbm_gbs"),
                                     "note": "algorithmic bytes (56 B/receiver + 29 B/heard pair + 32 B/tx) / sweep time: HBM is not the binding roofline"}},
                "verified": verified}
        if G == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(dep, lists[0], args)
        print(json.dumps(line), flush=True)
    med.close()
    if G > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------------
def run_c5(args):
    """BASELINE configs[4]: 1M nodes on 16 channels (node i on 11 + i % 16), every node moves by (U-0.5, U-0.5) m per
    tick (clamped to the area), 10k fresh sources per tick.  `value`: the tick alone, positions and tx records of the
    tick already in HBM (the dense position upload and the record gather happen before the start event);
    `e2e`: position upload from (pinned) host arrays + tick with host packet lists in and host result lists out."""
    import torch
    import torch.distributed as dist
    import radio_sim_b200 as rsb
    from radio_sim_b200.sharding import shard_bounds, TickPlan, merge_results
    import workloads as W

    rank, G, local, dev = setup_dist()
    dep = W.deployment("c5")
    dep["channels_used"] = 16
    N, T = dep["n"], dep["t"]
    # the position arrays live in pinned host memory (as the contract's e2e asks): the per-tick dense upload then
    # runs at PCIe speed instead of the driver's pageable staging
    px = torch.empty(N, dtype=torch.float64).pin_memory().numpy()
    py = torch.empty(N, dtype=torch.float64).pin_memory().numpy()
    px[:], py[:] = dep["x"], dep["y"]
    ch = dep["channel"]
    bounds = shard_bounds(N, G)
    lo, hi = bounds[rank]
    med = rsb.Medium("sinr", device=local, rx_base=lo, n_shards=G, compact=True)
    med.set_nodes(px[lo:hi], py[lo:hi], channel=ch[lo:hi])
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    med.set_stream(stream.cuda_stream)
    max_slots = int(T / G * 1.3) + 64
    if G > 1:
        open_peers(med, G, rank, max_slots, dev)
    rec_local = torch.zeros((max_slots if G > 1 else T) * 48, dtype=torch.uint8, device=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tick_no = [0]

    def next_tick():
        """positions and sources of the next tick (every rank draws the same java.util.Random streams)"""
        tick_no[0] += 1
        src = W.c5_tick(dep, tick_no[0])
        px[:], py[:] = dep["x"], dep["y"]
        return src, TickPlan(src, bounds)

    def merged_digest(res, plan):
        part = dict(tx=res.tx_column(), rx=res.rx.copy(), outcome=res.outcome.copy())
        if G > 1:
            parts = [None] * G
            dist.all_gather_object(parts, part)
            m = merge_results(plan, parts)
        else:
            m = part
        return digest(m["tx"], m["rx"], m["outcome"]), len(m["rx"])

    def tick_e2e(src, plan):
        med.update_positions(px[lo:hi], py[lo:hi])
        if G == 1:
            return med.tick(src, copy=False)
        return med.peer_tick(plan.local_src[rank], copy=False)

    # ticks 1..3 against the oracle's digests (tests/golden/fullsize_hashes.json)
    verified = None
    if not args.no_verify:
        ok = []
        for k in (1, 2, 3):
            src, plan = next_tick()
            d, n = merged_digest(tick_e2e(src, plan), plan)
            g = golden(f"c5-tick{k}")
            ok.append(bool(g and g["sha256"] == d and g["n_pairs"] == n))
        verified = {"golden": all(ok), "what": "tx, rx, outcome lists of mobility ticks 1-3"
                    + (" merged over the ranks" if G > 1 else "") + " vs tests/golden/fullsize_hashes.json (the oracle's lists)"}

    def timed(kind, steps, sampler=None):
        times, dec, pairs, sweep_ms, launches = [], 0, 0, [], 0
        for _ in range(steps):
            src, plan = next_tick()
            if kind != "e2e":
                med.update_positions(px[lo:hi], py[lo:hi])
                med.build_tx_records(plan.local_src[rank] if G > 1 else src, rec_local.data_ptr())
            torch.cuda.synchronize()
            if G > 1:
                dist.barrier()
                torch.cuda.synchronize()
            if sampler:
                sampler.active = True
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flush_buf.fill_(1)
            if G > 1:
                med.peer_barrier_enqueue()          # device-side rendezvous (see run_static)
            if kind == "e2e":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            if kind != "e2e":
                e0.record(stream)
                if G == 1:
                    med.tick_enqueue(T, rec_local.data_ptr())
                else:
                    med.peer_exchange_tick_enqueue(plan.tpad, rec_local.data_ptr())
                res = None
                e1.record(stream)
                torch.cuda.synchronize()
                wall = time.perf_counter() - t0
            else:
                res = tick_e2e(src, plan)       # synchronous: the lists are in host memory when it returns
                wall = time.perf_counter() - t0
            if sampler:
                sampler.active = False
            if res is None:
                res = med.tick_finish()
                times.append(e0.elapsed_time(e1) * 1e-3)
            else:
                times.append(wall)
            tm = med.timing()
            dec += res.n_decisions; pairs += res.n_pairs
            sweep_ms.append(tm["ms_sweep"]); launches += tm["kernel_launches"]
        return times, dec, pairs, sweep_ms, launches

    def pipelined(steps, sets=4):
        """The same host-lists-in / host-lists-out tick, K ticks timed as a whole, with the asynchronous forms of the
        C-ABI: tick k is enqueued, the positions of tick k+1 are uploaded beside it (rmb_update_positions_async), then
        tick k's lists are fetched.  Positions and transmitters cycle through `sets` pre-drawn consecutive ticks held
        in pinned host memory (drawing 2 M uniforms per tick on the host would dwarf the tick)."""
        pxs = [torch.empty(hi - lo, dtype=torch.float64).pin_memory().numpy() for _ in range(sets)]
        pys = [torch.empty(hi - lo, dtype=torch.float64).pin_memory().numpy() for _ in range(sets)]
        srcs, plans = [], []
        for j in range(sets):
            src, plan = next_tick()
            pxs[j][:], pys[j][:] = px[lo:hi], py[lo:hi]
            srcs.append(src); plans.append(plan)
        walls = []
        for phase_steps in (min(8, steps), steps):                   # warm-up pass, timed pass
            med.update_positions(pxs[0], pys[0])
            torch.cuda.synchronize()
            if G > 1:
                med.peer_barrier_enqueue()
                torch.cuda.synchronize()
            pairs = 0
            t0 = time.perf_counter()
            for k in range(phase_steps):
                j = k % sets
                if G == 1:
                    med.build_tx_records(srcs[j], rec_local.data_ptr())
                    med.tick_enqueue(T, rec_local.data_ptr())
                else:
                    med.peer_push(plans[j].local_src[rank])
                    med.tick_peers_enqueue()
                med.update_positions_async(pxs[(j + 1) % sets], pys[(j + 1) % sets])
                med.tick_finish()
                pairs += med.fetch(copy=False).n_pairs
            torch.cuda.synchronize()
            walls.append((time.perf_counter() - t0) / phase_steps)
        return walls[-1], pairs

    sampler = ClockSampler(local, period=0.0003)        # the timed windows are ~0.3 ms
    sampler.start()
    timed("resident", args.warmup)
    times, dec, _, sweep_st, launches = timed("resident", args.steps, sampler)
    med.set_tuning(1 << 24, 0)
    kev_steps = min(args.steps, 50)
    timed("resident", 2)
    ktimes, kdec, _, sweep_ev, _ = timed("resident", kev_steps)
    med.set_tuning(0, 0)
    timed("e2e", max(1, args.warmup // 2))
    e_times, e_dec, e_pairs, _, _ = timed("e2e", args.steps)
    p_wall, _ = pipelined(min(args.steps, 200))
    sampler.stop_flag = True
    sampler.join(timeout=2)
    ptot = torch.tensor([p_wall], dtype=torch.float64, device=dev)
    if G > 1:
        dist.all_reduce(ptot, op=dist.ReduceOp.MAX)
    p_wall = float(ptot[0])
    tot = torch.tensor([sum(times), sum(e_times)], dtype=torch.float64, device=dev)
    cnt = torch.tensor([float(dec), float(e_dec), float(e_pairs)], dtype=torch.float64, device=dev)
    if G > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    if rank == 0:
        clocks = sampler.summary()
        f_max = (clocks["sm_max_mhz"] or 1965.0) * 1e6
        ms_sweep = float(np.mean(sweep_ev))
        achieved = kdec / kev_steps / (ms_sweep * 1e-3)
        peak = SMS * XU_LANES * f_max
        line = {"metric": METRIC, "value": float(cnt[0]) / float(tot[0]), "unit": UNIT, "n_gpus": G, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * float(tot[0]) / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
                "config": config_of(args, dep, channels=16, parallelism=f"receivers sharded x{G}",
                                    l2="flushed (256 MiB write) before every timed step",
                                    timing="value: CUDA events around the enqueued tick (positions and records of the tick "
                                           "resident); e2e: host wall clock around position upload + tick with host lists"),
                "clocks": clocks,
                "e2e": {"value": float(cnt[1]) / float(tot[1]), "unit": UNIT, "ms_per_step": 1e3 * float(tot[1]) / args.steps,
                        "h2d_bytes_per_step": int(16 * (hi - lo) + 4 * T / G), "d2h_bytes_per_step": int(9 * float(cnt[2]) / args.steps / G + 5 * T),
                        "heard_pairs_per_tick": float(cnt[2]) / args.steps,
                        "pipelined": {"ms_per_step": 1e3 * p_wall, "value": float(cnt[1]) / args.steps / p_wall, "unit": UNIT,
                                      "what": "the same host-lists-in / host-lists-out tick through the asynchronous C-ABI forms, K ticks "
                                              "timed as a whole: tick k enqueued, positions of tick k+1 uploaded beside it "
                                              "(rmb_update_positions_async), tick k's lists fetched; max over ranks"}},
                "gpu_launches": int(launches),
                "roofline": {"bound": "sfu", "kernel": "sweep_kernel", "achieved": achieved / 1e9, "peak": peak / 1e9,
                             "unit": "G MUFU-op/s (1 rsqrt per decision; peak = 148 SM x 16 lanes x sm_max_mhz)",
                             "frac": achieved / peak, "ms_per_launch": ms_sweep, "ms_per_launch_device_stamps": float(np.mean(sweep_st)),
                             "traffic": traffic_for("c5", G), "measured_pipe_peaks": mufu_microbench()},
                "verified": verified}
        if G == 1 and not args.no_cpu:
            src = W.c5_tick(dep, tick_no[0] + 1)
            line["cpu_baseline"] = cpu_baseline(dep, src, args)
        print(json.dumps(line), flush=True)
    med.close()
    if G > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------------
C1_PAYLOAD = "0102030405"


def c1_session(factory):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from replay import Session, grid10
    s = Session(factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    return s


def run_c1(args):
    """BASELINE configs[0]: the reference's own CPU-runnable case.  One tick = one node broadcasts (ten ticks = every
    node once), driven through the JSON radio protocol (test/TestEmulator.java:91-125's command sequence: node-config-set,
    time-set, transmit, time-step replies) into the host mirror and the C-ABI.  Launch-bound: reported as latency."""
    import torch
    import radio_sim_b200 as rsb
    from radio_sim_b200.simulator import CudaRadioMedium
    import workloads as W
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the medium has no CPU fallback")
    if int(os.environ.get("RANK", "0")) != 0:
        return
    x, y = W.grid_c1()
    # (a) the tick alone through the C-ABI: rmb_tick of one packet on the 10-node UDGM medium
    med = rsb.Medium("udgm", compact=True)
    med.set_nodes(x, y)
    lat, dec = [], 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        r = med.tick(np.array([i % 10], dtype=np.int32), copy=False)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            lat.append(dt); dec += r.n_decisions
    tm = med.timing()
    med.close()
    # (b) the same through the JSON protocol and the adapter
    s = c1_session(lambda o, p: CudaRadioMedium("udgm"))
    emu = s.emus[0]
    plat, received = [], 0
    for i in range(args.warmup + args.steps):
        mark = len(emu.sent)
        t0 = time.perf_counter()
        s.step_to(1000 * (i + 1), [(emu, i % 10 + 1, 1000 * i, C1_PAYLOAD, {})])
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            plat.append(dt)
            received += sum(1 for m in emu.sent[mark:] if m.get("command") == "receive")
    s.sim.radio_medium.close()
    dep = dict(n=10, t=1)
    cpu = c1_cpu(args.steps)
    line = {"metric": METRIC, "value": dec / sum(lat), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(lat) / len(lat), "us_per_tick": 1e6 * float(np.median(lat)), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(args, dep, medium="udgm (UDGMRadioMedium defaults)", timing="host wall clock per tick (launch-bound: "
                                "latency, not throughput)"),
            "e2e": {"value": dec / sum(plat), "unit": UNIT, "us_per_tick": 1e6 * float(np.median(plat)),
                    "what": "time-set + transmit + time-step replies over the JSON protocol (in-memory harness, Python host mirror) "
                            "-> adapter -> rmb_tick -> receive messages", "receive_messages": received,
                    "h2d_bytes_per_step": 4 + 16, "d2h_bytes_per_step": 9 * 4 + 8},
            "gpu_launches": int(tm["kernel_launches"] * args.steps),
            "roofline": {"bound": "launch latency (9 decisions per tick: no throughput claim)", "kernel": "refmedia_count_kernel",
                         "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None},
            "cpu_baseline": cpu}
    print(json.dumps(line), flush=True)


def c1_cpu(steps):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc
    import workloads as W
    x, y = W.grid_c1()
    nodes, p = orc.NodeSet(x, y), orc.params(orc.MODEL_UDGM)
    lat, dec = [], 0
    for i in range(steps + 5):
        pk = orc.make_packets(np.array([i % 10], dtype=np.int32), 0.0, 26)
        t0 = time.perf_counter()
        r = orc.tick(p, nodes, pk)
        dt = time.perf_counter() - t0
        if i >= 5:
            lat.append(dt); dec += r["n_decisions"]
    return {"value": dec / sum(lat), "unit": UNIT, "cores": 1, "kind": "port", "us_per_tick": 1e6 * float(np.median(lat)),
            "sample": "the whole tick (10 nodes x 1 packet), through ctypes",
            "single_thread": {"value": dec / sum(lat), "unit": UNIT, "cores": 1}}


def run_c1_reference(args):
    cpu = c1_cpu(max(args.steps, 10))
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": cpu["us_per_tick"] / 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_of(args, dict(n=10, t=1)),
            "cpu_baseline": cpu, "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
def run_single_process(args):
    """N GPUs behind ONE object in ONE process (rmb_group_*: what one JVM with the adapter does): a tick is one call
    with the whole packet list, inside the library one host worker thread per GPU pushes its share, runs its shard's
    tick and brings its lists to host memory; the call returns when every shard's lists are there (rank order == node
    order, so packet t's rows are the shards' rows of t one after the other -- the merge is a walk over tx_first, no
    copy).  Reported as e2e."""
    import torch
    import radio_sim_b200 as rsb
    import workloads as W
    G = args.gpus
    if torch.cuda.device_count() < G:
        raise SystemExit(f"--single-process --gpus {G} needs {G} visible GPUs")
    name = args.workload
    dep = W.deployment("c4" if name == "c4-grid" else name, grid=name == "c4-grid")
    N, T = dep["n"], dep["t"]
    lists = source_lists(dep, max(1, min(args.steps, 8)))
    grp = rsb.Group("sinr", devices=tuple(range(G)), max_tx_per_tick=T, compact=True, **medium_params(args))
    grp.set_nodes(dep["x"], dep["y"])
    walls, dec, pairs = [], 0, 0
    for i in range(args.warmup + args.steps):
        for r in range(G):
            torch.cuda.synchronize(r)
        t0 = time.perf_counter()
        res = grp.tick(lists[i % len(lists)], copy=False)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            walls.append(dt)
            dec += res.n_decisions; pairs += res.n_pairs
    m = grp.tick(lists[0]).merged()
    d = digest(m["tx"], m["rx"], m["outcome"])
    gold = golden(name)
    grp.close()
    line = {"metric": METRIC, "value": dec / sum(walls), "unit": UNIT, "n_gpus": G, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(walls) / len(walls), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32+f64", "data": "synthetic",
            "config": config_of(args, dep, parallelism=f"receivers sharded x{G}, ONE process (rmb_group_*), one host worker thread per GPU inside the library",
                                timing="host wall clock around rmb_group_tick: host packet list in, every shard's lists in host memory out"),
            "e2e": {"value": dec / sum(walls), "unit": UNIT, "ms_per_step": 1e3 * sum(walls) / len(walls),
                    "h2d_bytes_per_step": int(4 * T), "d2h_bytes_per_step": int(9 * pairs / len(walls) + 5 * T),
                    "heard_pairs": int(pairs / len(walls))},
            "gpu_launches": 3 * G * args.steps,
            "verified": {"sha256": d, "golden": None if gold is None else bool(gold["sha256"] == d), "heard_pairs": int(len(m["rx"])),
                         "what": "the shards' lists merged in the reference's order vs tests/golden/fullsize_hashes.json"}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=["c1", "c2", "c3", "c4", "c4-grid", "c5"],
                    help="c4 = BASELINE configs[3] (the metric's workload, default)")
    ap.add_argument("--grid", action="store_true", help="same as --workload c4-grid")
    ap.add_argument("--pl-exponent", type=float, default=None, help="path-loss exponent (default 3.0; 2.7 = general exponent, 2 MUFU per decision)")
    ap.add_argument("--tx-power", type=float, default=None, help="tx power of every node in dBm (CC2420 levels: -25, -15, -7, 0)")
    ap.add_argument("--rx-prob", type=float, default=None, help="rxProbability of every node (< 1: stochastic reception, N = 1)")
    ap.add_argument("--variant", type=lambda v: int(v, 0), default=0, help="rmb_set_tuning word (kernel variant, splits, ...)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-verify", action="store_true", help="skip the digest check of the result lists")
    ap.add_argument("--single-process", action="store_true", help="N handles in one process, one host thread per GPU")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="N>1: how the tick's tx records travel")
    args = ap.parse_args()
    if args.grid:
        args.workload = "c4-grid"
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    elif args.single_process:
        run_single_process(args)
    elif args.workload == "c1":
        run_c1(args)
    elif args.workload == "c5":
        run_c5(args)
    else:
        run_static(args)


if __name__ == "__main__":
    main()
