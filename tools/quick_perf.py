"""Ad-hoc sweep-variant timing on one GPU (development aid; bench.py is the contract)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import radio_sim_b200 as rsb

def run(n, t, side, variants, channels=None, reps=3):
    rng = np.random.default_rng(3)
    x, y = rng.uniform(0, side, n), rng.uniform(0, side, n)
    ch = None if channels is None else (11 + np.arange(n) % channels).astype(np.int32)
    src = rng.permutation(n)[:t].astype(np.int32)
    with rsb.Medium("sinr") as m:
        m.set_nodes(x, y, channel=ch)
        for v in variants:
            m.set_tuning(v, 0)
            best = None
            for _ in range(reps):
                t0 = time.perf_counter(); r = m.tick(src, copy=False); wall = time.perf_counter() - t0
                tm = m.timing()
                if best is None or tm["ms_sweep"] < best[0]["ms_sweep"]:
                    best = (tm, wall)
            tm, wall = best
            print(f"n={n} t={t} ch={channels} variant={v} pairs={r.n_pairs} dec={r.n_decisions:.3e} recheck={r.n_recheck} K={r.near_slots} "
                  f"sweep={tm['ms_sweep']:.3f}ms resolve={tm['ms_resolve']:.3f} emit={tm['ms_emit']:.3f} prep={tm['ms_prepare']:.3f} "
                  f"total={tm['ms_total']:.3f} wall={wall*1e3:.2f}ms sweep_rate={r.n_decisions/tm['ms_sweep']/1e9*1e3:.1f} Gpairs/s", flush=True)

if __name__ == "__main__":
    v = [0, 2, 3, 4, 5]
    run(1000, 50, 1000.0, [0])
    run(100_000, 1000, 3162.0, v)
    run(1_000_000, 10_000, 10_000.0, v, reps=2)
    run(1_000_000, 10_000, 10_000.0, [0], channels=16, reps=2)
