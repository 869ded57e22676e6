"""The synthetic inputs of BASELINE.json's configs, all drawn from java.util.Random so that a
Java harness regenerates them bit for bit (radio-sim_b200/java/tools/GenGolden.java uses the same loops).

  C1  10 nodes at (30 i, 30 j), i in 0..4, j in 0..1; every node broadcasts once (UDGM defaults)
  C2  seed 1: 1 000 nodes uniform in [0, 1000)^2, 50 transmitters
  C3  seed 2: 100 000 nodes uniform in [0, 3162)^2, 1 000 transmitters
  C4  seed 3: 1 000 000 nodes uniform in [0, 10 000)^2, 10 000 transmitters;  grid variant: 1000 x 1000, 10 m pitch
  C5  seed 4: C4 placement, node i on channel 11 + (i mod 16); tick k >= 1 draws from Random(4 + k): every node
      moves by (U - 0.5, U - 0.5) m clamped to the area, then 10 000 fresh transmitters

Placement loop: for i in 0..n-1 { x[i] = nextDouble() * side; y[i] = nextDouble() * side; }  z = 0.
Transmitters: the first T entries of a Fisher-Yates shuffle of 0..n-1 continued from the same generator
(JavaRandom.sample_distinct).  Used by bench.py, tests/ and tools/.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from radio_sim_b200.javarandom import JavaRandom   # noqa: E402

CONFIGS = {
    "c2": dict(seed=1, n=1_000, side=1000.0, t=50),
    "c3": dict(seed=2, n=100_000, side=3162.0, t=1_000),
    "c4": dict(seed=3, n=1_000_000, side=10_000.0, t=10_000),
    "c5": dict(seed=4, n=1_000_000, side=10_000.0, t=10_000, channels=16),
}


def grid_c1():
    """BASELINE configs[0]: 5 x 2 grid, 30 m pitch; node i broadcasts in tick i."""
    x = np.array([30.0 * (i % 5) for i in range(10)])
    y = np.array([30.0 * (i // 5) for i in range(10)])
    return x, y


def random_placement(rng, n, side):
    u = rng.next_doubles(2 * n)
    return u[0::2] * side, u[1::2] * side


def deployment(name, grid=False, n=None, t=None, side=None):
    """-> dict(x, y, channel, src, side, n, t).  n/t/side override the config's (scaled-down test cases)."""
    c = dict(CONFIGS[name])
    if n is not None:
        c["n"] = n
    if t is not None:
        c["t"] = t
    if side is not None:
        c["side"] = side
    rng = JavaRandom(c["seed"])
    if grid:
        k = int(round(c["n"] ** 0.5))
        assert k * k == c["n"], "grid deployment needs a square node count"
        pitch = c["side"] / k
        i = np.arange(c["n"])
        x, y = (i % k) * pitch, (i // k) * pitch
    else:
        x, y = random_placement(rng, c["n"], c["side"])
    ch = np.full(c["n"], 26, dtype=np.int32)
    if c.get("channels"):
        ch = (11 + np.arange(c["n"]) % c["channels"]).astype(np.int32)
    src = rng.sample_distinct(c["n"], c["t"])
    return dict(x=np.ascontiguousarray(x, dtype=np.float64), y=np.ascontiguousarray(y, dtype=np.float64), channel=ch,
                src=src, side=c["side"], n=c["n"], t=c["t"], seed=c["seed"])


def c5_tick(dep, k):
    """Mobility tick k >= 1 of C5, in place: moves every node, returns the tick's transmitters."""
    rng = JavaRandom(dep["seed"] + k)
    u = rng.next_doubles(2 * dep["n"])
    hi = np.nextafter(dep["side"], 0.0)
    dep["x"] = np.clip(dep["x"] + (u[0::2] - 0.5), 0.0, hi)
    dep["y"] = np.clip(dep["y"] + (u[1::2] - 0.5), 0.0, hi)
    return rng.sample_distinct(dep["n"], dep["t"])
