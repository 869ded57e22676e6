"""Per-phase device times of one tick (development aid).
usage: phases.py [n] [t] [side] [channels] [tuning,tuning,...]   tuning = variant | splits << 8"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, time
import radio_sim_b200 as rsb
a = sys.argv
n, t, side = int(a[1]) if len(a) > 1 else 1_000_000, int(a[2]) if len(a) > 2 else 10_000, float(a[3]) if len(a) > 3 else 10_000.0
ch_n = int(a[4]) if len(a) > 4 else 0
tunings = [int(v, 0) for v in a[5].split(',')] if len(a) > 5 else [0]
rng = np.random.default_rng(3)
x, y = rng.uniform(0, side, n), rng.uniform(0, side, n)
src = rng.permutation(n)[:t].astype(np.int32)
ch = (11 + np.arange(n) % ch_n).astype(np.int32) if ch_n else None
with rsb.Medium("sinr") as m:
    m.set_nodes(x, y, channel=ch)
    for tune in tunings:
        m.set_tuning(tune, 0)
        for i in range(6):
            t0 = time.perf_counter(); r = m.tick(src, copy=False); wall = time.perf_counter() - t0
            tm = m.timing()
        print("tune", hex(tune), {k: round(v, 4) if isinstance(v, float) else v for k, v in tm.items()},
              "wall_ms", round(wall * 1e3, 3), "pairs", r.n_pairs, "recheck", r.n_recheck)
