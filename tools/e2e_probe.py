"""Per-step host wall time of rmb_tick (host lists in, host lists out) on C4, one GPU (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np, torch
import radio_sim_b200 as rsb
import workloads as W
import bench
dep = W.deployment("c4")
lists = bench.source_lists(dep, int(sys.argv[1]) if len(sys.argv) > 1 else 8)
med = rsb.Medium("sinr", compact=True)
med.set_nodes(dep["x"], dep["y"])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = []
for i in range(24):
    flush.fill_(1); torch.cuda.synchronize()
    t0 = time.perf_counter(); r = med.tick(lists[i % len(lists)], copy=False); dt = time.perf_counter() - t0
    tm = med.timing()
    out.append((round(dt * 1e3, 3), r.n_pairs, round(tm["ms_total"], 3), r.near_slots))
print(out)
med.close()
