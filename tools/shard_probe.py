"""One rank's share of the G-way sharded C4 tick, on ONE GPU (development aid): where do the
fixed per-tick costs go once the sweep is 1/G of the work?
usage: shard_probe.py [G] [reps] [tuning]     (run under ncu for the per-kernel list)
tuning bit 24 (0x1000000): CUDA events between the tick's kernels (serialised launches) instead of the kernels' own
time stamps (programmatic dependent launch active)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import radio_sim_b200 as rsb
import workloads as W

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
tuning = int(sys.argv[3], 0) if len(sys.argv) > 3 else 0
dep = W.deployment("c4")
N, T, x, y, src = dep["n"], dep["t"], dep["x"], dep["y"], dep["src"]
dev = torch.device("cuda", 0)
rec = torch.zeros(T * 48, dtype=torch.uint8, device=dev)
with rsb.Medium("sinr") as full:
    full.set_nodes(x, y)
    full.build_tx_records(src, rec.data_ptr())
torch.cuda.synchronize()
lo, hi = 0, N // G
med = rsb.Medium("sinr", rx_base=lo, n_shards=G)
med.set_nodes(x[lo:hi], y[lo:hi])
med.set_tuning(tuning, 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream); med.set_stream(stream.cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
rows = []
for i in range(reps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.fill_(1)                                   # the host runs ahead while the flush executes: no launch latency in the timed region
    t0 = time.perf_counter(); e0.record(stream); med.tick_enqueue(T, rec.data_ptr()); e1.record(stream)
    torch.cuda.synchronize(); wall = time.perf_counter() - t0
    r = med.tick_finish()
    tm = med.timing()
    rows.append((e0.elapsed_time(e1) * 1e3, wall * 1e6, tm["ms_prepare"] * 1e3, tm["ms_sweep"] * 1e3, tm["ms_resolve"] * 1e3,
                 tm["ms_emit"] * 1e3, tm["ms_total"] * 1e3))
rows = np.array(rows[min(3, reps - 1):])
print("G", G, "tuning", hex(tuning), "pairs", r.n_pairs, "decisions", r.n_decisions, "launches", tm["kernel_launches"])
print("median us: events %.1f wall %.1f | prepare %.1f sweep %.1f resolve %.1f emit %.1f total %.1f" % tuple(np.median(rows, axis=0)))
med.close()
