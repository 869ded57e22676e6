"""One rank's share of the G-way sharded C5 tick (16 channels, mobility) on ONE GPU: device phases and the host-side
cost of the per-tick calls.   usage: c5_shard_probe.py [G] [ticks]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import radio_sim_b200 as rsb
import workloads as W

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 24
dep = W.deployment("c5")
N, T = dep["n"], dep["t"]
dev = torch.device("cuda", 0)
lo, hi = 0, N // G
px = torch.empty(N, dtype=torch.float64).pin_memory().numpy(); py = torch.empty(N, dtype=torch.float64).pin_memory().numpy()
full = rsb.Medium("sinr"); full.set_nodes(dep["x"], dep["y"], channel=dep["channel"])          # builds every rank's records
med = rsb.Medium("sinr", rx_base=lo, n_shards=G, compact=True)
med.set_nodes(dep["x"][lo:hi], dep["y"][lo:hi], channel=dep["channel"][lo:hi])
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream); med.set_stream(stream.cuda_stream)
rec = torch.zeros(T * 48, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for mode in ((0, 1 << 24) if len(sys.argv) <= 3 else [int(v, 0) for v in sys.argv[3].split(",")]):
    med.set_tuning(mode, 0)
    rows = []
    for k in range(1, ticks + 1):
        src = W.c5_tick(dep, k if mode == 0 else ticks + k)
        px[:], py[:] = dep["x"], dep["y"]
        full.update_positions(px, py); full.build_tx_records(src, rec.data_ptr()); torch.cuda.synchronize()
        t0 = time.perf_counter(); med.update_positions(px[lo:hi], py[lo:hi]); t_up = time.perf_counter() - t0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flush.fill_(1)
        e0.record(stream); med.tick_enqueue(T, rec.data_ptr()); e1.record(stream); torch.cuda.synchronize()
        r = med.tick_finish(); tm = med.timing()
        t0 = time.perf_counter(); rr = med.tick_records(T, rec.data_ptr(), copy=False); t_tick = time.perf_counter() - t0
        rows.append((e0.elapsed_time(e1) * 1e3, tm["ms_prepare"] * 1e3, tm["ms_sweep"] * 1e3, tm["ms_resolve"] * 1e3, tm["ms_emit"] * 1e3,
                     t_up * 1e6, t_tick * 1e6))
    rows = np.array(rows[4:])
    print("G", G, "mode", hex(mode), "pairs", r.n_pairs, "decisions", r.n_decisions)
    print("median us: tick (events) %.1f | prepare %.1f sweep %.1f resolve %.1f emit %.1f || host: update_positions %.1f, tick_records (sync, D2H) %.1f"
          % tuple(np.median(rows, axis=0)))
    print("update_positions by tick mod 8:", [round(float(np.median(rows[(np.arange(len(rows)) + 4) % 8 == j, 5])), 0) for j in range(8)])
med.close(); full.close()
