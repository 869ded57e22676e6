"""Where does a C5 tick through the host API spend its time? (development aid)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import radio_sim_b200 as rsb
N, T = 1_000_000, 10_000
rng = np.random.default_rng(4)
x = torch.empty(N, dtype=torch.float64).pin_memory().numpy(); y = torch.empty(N, dtype=torch.float64).pin_memory().numpy()
x[:], y[:] = rng.uniform(0, 1e4, N), rng.uniform(0, 1e4, N)
ch = (11 + np.arange(N) % 16).astype(np.int32)
src = rng.permutation(N)[:T].astype(np.int32)
with rsb.Medium("sinr", compact=True) as m:
    m.set_nodes(x, y, channel=ch)
    for name, fn in [("update_positions dense (pinned)", lambda: m.update_positions(x, y)),
                     ("tick host lists", lambda: m.tick(src, copy=False)),
                     ("update + tick", lambda: (m.update_positions(x, y), m.tick(src, copy=False)))]:
        ts = []
        for i in range(12):
            x += 0.1; torch.cuda.synchronize()
            t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        print(f"{name}: {np.median(ts[2:]):.3f} ms")
