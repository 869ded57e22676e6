"""Per-phase device times of one C4 tick (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, time
import radio_sim_b200 as rsb
n, t, side = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 10_000, float(sys.argv[3]) if len(sys.argv) > 3 else 10_000.0
ch_n = int(sys.argv[4]) if len(sys.argv) > 4 else 0
rng = np.random.default_rng(3)
x, y = rng.uniform(0, side, n), rng.uniform(0, side, n)
src = rng.permutation(n)[:t].astype(np.int32)
ch = (11 + np.arange(n) % ch_n).astype(np.int32) if ch_n else None
with rsb.Medium("sinr") as m:
    m.set_nodes(x, y, channel=ch)
    for i in range(6):
        t0 = time.perf_counter(); r = m.tick(src, copy=False); wall = time.perf_counter() - t0
        tm = m.timing()
    print({k: round(v, 4) if isinstance(v, float) else v for k, v in tm.items()}, "wall_ms", round(wall * 1e3, 3), "pairs", r.n_pairs, "recheck", r.n_recheck)
