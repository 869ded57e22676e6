"""The C5 tick (1M nodes on 16 channels, 10k sources) with the records resident, per phase, for a list of tuning words.
usage: c5_tick_probe.py [tuning ...]   (development aid)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import radio_sim_b200 as rsb
N, T = 1_000_000, 10_000
rng = np.random.default_rng(4)
x, y = rng.uniform(0, 1e4, N), rng.uniform(0, 1e4, N)
ch = (11 + np.arange(N) % 16).astype(np.int32)
src = rng.permutation(N)[:T].astype(np.int32)
dev = torch.device("cuda", 0)
rec = torch.zeros(T * 48, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for word in [int(a, 0) for a in sys.argv[1:]] or [0]:
    with rsb.Medium("sinr", compact=True) as m:
        m.set_nodes(x, y, channel=ch)
        m.set_tuning(word, 0)
        m.build_tx_records(src, rec.data_ptr())
        rows = []
        for i in range(16):
            flush.fill_(1); torch.cuda.synchronize()
            r = m.tick_resident(T, rec.data_ptr())
            tm = m.timing()
            rows.append((tm["ms_prepare"] * 1e3, tm["ms_sweep"] * 1e3, tm["ms_emit"] * 1e3, tm["ms_total"] * 1e3))
        print("tuning %#x pairs %d | prepare %.1f sweep %.1f emit %.1f total %.1f us" % ((word, r.n_pairs) + tuple(np.median(np.array(rows[3:]), axis=0))))
