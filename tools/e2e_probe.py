"""Per-step host wall time of the synchronous tick (lists out to host memory) on one GPU (development aid).
usage: e2e_probe.py [workload c2|c3|c4] [G]
G = 1: rmb_tick, host packet list in, host lists out.  G > 1: one rank's share of the G-way sharded tick through
rmb_tick_records (all records resident, this shard's receivers): what is left of a tick's host side once the sweep is
1/G of the work.  RMB_SLAB_GUESS=0 switches the guessed copy behind the finish kernel off (second synchronisation back)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import radio_sim_b200 as rsb
import workloads as W
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dep = W.deployment(wl)
N, T, x, y = dep["n"], dep["t"], dep["x"], dep["y"]
lists = bench.source_lists(dep, 8)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
if G == 1:
    med = rsb.Medium("sinr", compact=True)
    med.set_nodes(x, y)
    call = lambda k: med.tick(lists[k], copy=False)
else:
    recs = []
    with rsb.Medium("sinr") as full:
        full.set_nodes(x, y)
        for src in lists:
            rec = torch.zeros(T * 48, dtype=torch.uint8, device="cuda")
            full.build_tx_records(src, rec.data_ptr())
            recs.append(rec)
    torch.cuda.synchronize()
    med = rsb.Medium("sinr", compact=True, rx_base=0, n_shards=G)
    med.set_nodes(x[:N // G], y[:N // G])
    call = lambda k: med.tick_records(T, recs[k].data_ptr(), copy=False)
wall, devt = [], []
for i in range(40):
    flush.fill_(1); torch.cuda.synchronize()
    t0 = time.perf_counter(); r = call(i % len(lists)); dt = time.perf_counter() - t0
    if i >= 8:
        wall.append(dt * 1e6); devt.append(med.timing()["ms_total"] * 1e3)
print("%s G %d guess %s: pairs %d | host-to-host median %.1f us (p10 %.1f, p90 %.1f) | device %.1f us | host side %.1f us"
      % (wl, G, os.environ.get("RMB_SLAB_GUESS", "1"), r.n_pairs, np.median(wall), np.percentile(wall, 10), np.percentile(wall, 90),
         np.median(devt), np.median(wall) - np.median(devt)))
med.close()
