"""CPU: size-independent properties of the oracle's SINR medium (the checker itself must be self-consistent
before the CUDA path is held against it): invariance under node renumbering, independence of other channels,
the three-way outcome in terms of the reported values, the decision count, thread-count independence, and the
interference sum being monotone in the set of transmitters."""
import numpy as np
import pytest

from helpers import random_deployment, pick_sources


def _tick(orc, x, y, z, ch, src, power=0.0, pkt_ch=None, **kw):
    pc = ch[src] if pkt_ch is None else pkt_ch
    return orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch, **kw), orc.make_packets(src, power, pc),
                    want_total=True)


def test_renumbering_the_nodes_only_renumbers_the_lists(orc):
    n, t = 1500, 60
    x, y, z, ch = random_deployment(71, n, 500.0, channels=3)
    src = pick_sources(71, n, t)
    a = _tick(orc, x, y, z, ch, src)
    perm = np.random.default_rng(71).permutation(n)          # new index i holds old node perm[i]
    inv = np.empty(n, dtype=np.int64); inv[perm] = np.arange(n)
    b = _tick(orc, x[perm], y[perm], z[perm], ch[perm], inv[src].astype(np.int32))
    pa, pb = a["pairs"], b["pairs"]
    order = np.lexsort((perm[pb["rx"]], pb["tx"]))           # b's rows in a's (packet, old node index) order
    assert len(pa) == len(pb) and np.array_equal(pa["tx"], pb["tx"][order])
    assert np.array_equal(perm[pb["rx"]][order], pa["rx"])
    assert np.array_equal(pb["outcome"][order], pa["outcome"])
    assert np.array_equal(pb["rssi_dbm"][order], pa["rssi_dbm"])      # a pair's link budget does not depend on the numbering
    assert np.array_equal(pb["sinr_db"][order], pa["sinr_db"])        # nor does the sum: packets keep their arrival order
    assert np.array_equal(b["total_dbm"], a["total_dbm"][perm])


def test_other_channels_are_invisible(orc):
    n = 1200
    x, y, z, ch = random_deployment(72, n, 400.0, channels=2)          # channels 11 and 12
    src = pick_sources(72, n, 80)
    on11 = src[ch[src] == ch.min()]
    both = _tick(orc, x, y, z, ch, src)
    only = _tick(orc, x, y, z, ch, on11)
    rows = both["pairs"][np.isin(both["pairs"]["tx"], np.flatnonzero(ch[src] == ch.min()))]
    assert len(rows) == len(only["pairs"])
    assert np.array_equal(rows["rx"], only["pairs"]["rx"]) and np.array_equal(rows["outcome"], only["pairs"]["outcome"])
    assert np.array_equal(rows["sinr_db"], only["pairs"]["sinr_db"])


def test_outcome_follows_from_the_reported_values(orc):
    n = 2500
    x, y, z, ch = random_deployment(73, n, 600.0)
    src = pick_sources(73, n, 150)
    p = orc.params(orc.MODEL_SINR)
    r = _tick(orc, x, y, z, ch, src)
    pr = r["pairs"]
    assert len(pr) > 1000
    assert (pr["rssi_dbm"] >= p.sensitivity_dbm).all()                 # only heard pairs are listed
    assert np.array_equal(pr["outcome"] == orc.DELIVERED, pr["sinr_db"] >= p.sinr_threshold_db)
    assert (pr["rx"] != src[pr["tx"]]).all()                           # nobody receives their own frame
    assert r["n_decisions"] == len(src) * (n - 1)                      # one channel, everybody enabled
    assert np.all(np.diff(pr["tx"]) >= 0)                              # (packet arrival, node index) order
    same = np.diff(pr["tx"]) == 0
    assert np.all(np.diff(pr["rx"])[same] > 0)


def test_more_transmitters_never_raise_a_sinr(orc):
    n = 1500
    x, y, z, ch = random_deployment(74, n, 450.0)
    src = pick_sources(74, n, 90)
    few = _tick(orc, x, y, z, ch, src[:30])
    many = _tick(orc, x, y, z, ch, src)
    head = many["pairs"][many["pairs"]["tx"] < 30]
    assert np.array_equal(head["tx"], few["pairs"]["tx"]) and np.array_equal(head["rx"], few["pairs"]["rx"])
    assert np.array_equal(head["rssi_dbm"], few["pairs"]["rssi_dbm"])
    assert (head["sinr_db"] <= few["pairs"]["sinr_db"]).all()
    assert (many["total_dbm"] >= few["total_dbm"]).all()


@pytest.mark.parametrize("threads", [2, 7])
def test_thread_count_does_not_change_a_bit(orc, threads):
    n = 1800
    x, y, z, ch = random_deployment(75, n, 500.0, channels=4)
    src = pick_sources(75, n, 70)
    p = orc.params(orc.MODEL_SINR)
    nodes, pk = orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src])
    a = orc.tick(p, nodes, pk, want_total=True, threads=1)
    b = orc.tick(p, nodes, pk, want_total=True, threads=threads)
    assert a["pairs"].tobytes() == b["pairs"].tobytes() and np.array_equal(a["total_dbm"], b["total_dbm"])
    assert a["n_decisions"] == b["n_decisions"]
