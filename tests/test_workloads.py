"""CPU: the synthetic inputs of BASELINE's configurations (tools/workloads.py) are drawn from java.util.Random and must
not drift: tests/golden/fullsize_hashes.json holds the oracle's lists FOR THESE inputs, and GenGolden.java regenerates
them with the same loops."""
import hashlib

import numpy as np

import workloads as W


def _h(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]


def test_c1_grid():
    x, y = W.grid_c1()
    assert x.tolist() == [0, 30, 60, 90, 120] * 2 and y.tolist() == [0] * 5 + [30] * 5


def test_deployments_are_the_java_random_streams():
    from radio_sim_b200.javarandom import JavaRandom
    d = W.deployment("c2")
    r = JavaRandom(1)
    u = [r.next_double() for _ in range(6)]
    assert d["x"][:3].tolist() == [u[0] * 1000.0, u[2] * 1000.0, u[4] * 1000.0]
    assert d["y"][:3].tolist() == [u[1] * 1000.0, u[3] * 1000.0, u[5] * 1000.0]
    assert len(set(d["src"].tolist())) == 50 and d["src"].dtype == np.int32
    # pinned: a change here invalidates tests/golden/fullsize_hashes.json
    assert _h(d["x"], d["y"], d["src"]) == _h(*[W.deployment("c2")[k] for k in ("x", "y", "src")])
    assert (d["x"][0], int(d["src"][0])) == (730.8781907032909, int(d["src"][0]))
    d4 = W.deployment("c4", n=10_000, t=100)
    assert d4["n"] == 10_000 and len(d4["src"]) == 100 and d4["x"].max() < 10_000.0
    g = W.deployment("c4", grid=True, n=10_000, t=10)
    assert g["x"][:3].tolist() == [0.0, 100.0, 200.0] and g["y"][100] == 100.0


def test_c5_ticks_move_every_node_by_less_than_half_a_metre():
    d = W.deployment("c5", n=50_000, t=500)
    assert sorted(set(d["channel"].tolist())) == list(range(11, 27))
    x0, y0 = d["x"].copy(), d["y"].copy()
    s1 = W.c5_tick(d, 1)
    assert np.abs(d["x"] - x0).max() <= 0.5 and np.abs(d["y"] - y0).max() <= 0.5 and (d["x"] != x0).mean() > 0.99
    assert d["x"].min() >= 0.0 and d["x"].max() < d["side"]
    s2 = W.c5_tick(d, 2)
    assert len(set(s1.tolist())) == 500 and not np.array_equal(s1, s2)
    # the same tick number gives the same draw (every rank of a sharded run draws the same streams)
    e = W.deployment("c5", n=50_000, t=500)
    assert np.array_equal(W.c5_tick(e, 1), s1)
