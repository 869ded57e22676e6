"""Sweep-variant x tx-split tuning for one rank's share of the G-way sharded C4 tick (development aid).
usage: shard_tune.py G variant,variant,... split,split,..."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import radio_sim_b200 as rsb
import bench

G = int(sys.argv[1]); variants = [int(v) for v in sys.argv[2].split(',')]; splits = [int(v) for v in sys.argv[3].split(',')]
N, T = bench.N_NODES, bench.N_TX
x, y, src = bench.deployment()
dev = torch.device("cuda", 0)
rec = torch.zeros(T * 48, dtype=torch.uint8, device=dev)
with rsb.Medium("sinr") as full:
    full.set_nodes(x, y); full.build_tx_records(src, rec.data_ptr())
torch.cuda.synchronize()
med = rsb.Medium("sinr", rx_base=0, n_shards=G)
med.set_nodes(x[:N // G], y[:N // G])
for v in variants:
    out = []
    for s in splits:
        med.set_tuning(v | (s << 8), 0)
        ts = []
        for i in range(8):
            med.tick_resident(T, rec.data_ptr()); ts.append(med.timing()["ms_sweep"] * 1e3)
        out.append(f"S={s}:{np.median(ts[2:]):.1f}")
    print(f"G={G} variant={v}  " + "  ".join(out), flush=True)
med.close()
