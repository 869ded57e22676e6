"""-m gpu: rmb_group_* -- several handles behind one object (one process, one worker thread per device inside the
library).  A one-GPU box runs it with the same device named more than once; the lists must equal the single handle's
and the oracle's, row for row, after the merge the header describes (packet t's rows = shard 0's rows of t, then
shard 1's, ...)."""
import numpy as np
import pytest
from helpers import random_deployment, pick_sources, compare, SINR_TOL_DB

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rsb():
    import radio_sim_b200
    return radio_sim_b200


class _Merged:
    """GroupResult.merged() in the shape helpers.compare expects"""

    def __init__(self, gr, f64=True):
        m = gr.merged()
        self.tx, self.rx, self.outcome = m["tx"], m["rx"], m["outcome"]
        self.rssi_dbm = m["rssi_dbm"]
        self.n_pairs, self.n_decisions, self.n_delivered = gr.n_pairs, gr.n_decisions, gr.n_delivered
        self.n_interfered = gr.n_pairs - gr.n_delivered


@pytest.mark.parametrize("devices", [(0,), (0, 0), (0, 0, 0)])
def test_group_equals_the_oracle(rsb, orc, devices):
    x, y, z, ch = random_deployment(71, 30_000, 1500.0, channels=3)
    src = pick_sources(71, 30_000, 400)
    power = np.where(np.arange(400) % 3 == 0, -7.0, 0.0)
    with rsb.Group("sinr", devices=devices, max_tx_per_tick=512, compact=True) as g:
        g.set_nodes(x, y, z, channel=ch)
        r = g.tick(src, power_dbm=power)
        rssi, recv = g.node_info_f64()
    ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, power, ch[src]),
                   threads=8, want_total=True)
    m = r.merged()
    p = ref["pairs"]
    assert r.n_pairs == ref["n_pairs"] > 1000 and r.n_decisions == ref["n_decisions"]
    assert np.array_equal(m["tx"], p["tx"]) and np.array_equal(m["rx"], p["rx"]) and np.array_equal(m["outcome"], p["outcome"])
    assert np.abs(m["rssi_dbm"].astype(np.float64) - p["rssi_dbm"]).max() < SINR_TOL_DB
    assert np.abs(rssi - ref["total_dbm"]).max() < SINR_TOL_DB
    heard = np.zeros(len(x), dtype=np.uint8); heard[p["rx"]] = 1
    assert np.array_equal(recv, heard)
    # the walk the header describes, without sorting: packet t's rows are shard 0's rows of t, then shard 1's, ...
    if len(devices) > 1:
        pos_of = np.full(len(src), -1); pos_of[r.packet_of[r.packet_of >= 0]] = np.flatnonzero(r.packet_of >= 0)
        rows = []
        for t in range(len(src)):
            for part in r.parts:
                lo, hi = part.tx_first[pos_of[t]], part.tx_first[pos_of[t] + 1]
                rows.extend((t, int(v)) for v in part.rx[lo:hi])
        assert rows == list(zip(p["tx"].tolist(), p["rx"].tolist()))


def test_group_updates_and_repartition(rsb, orc):
    x, y, z, ch = random_deployment(72, 20_000, 1200.0)
    src = pick_sources(72, 20_000, 300)
    rng = np.random.default_rng(73)
    with rsb.Group("sinr", devices=(0, 0), max_tx_per_tick=512, compact=True) as g:
        g.set_nodes(x, y, z, channel=ch)
        g.tick(src)
        # sparse moves across both shards, one radio switched off, one retuned
        idx = np.sort(rng.permutation(20_000)[:500]).astype(np.int32)
        x[idx] += rng.uniform(-30, 30, 500); y[idx] += rng.uniform(-30, 30, 500)
        g.update_positions(x[idx], y[idx], idx=idx)
        en = np.ones(20_000, dtype=np.uint8); en[[5, 15_000]] = 0
        ch2 = ch.copy(); ch2[[7, 12_345]] = 11
        g.update_radio(np.array([5, 15_000, 7, 12_345], dtype=np.int32), channel=ch2[[5, 15_000, 7, 12_345]],
                       enabled=en[[5, 15_000, 7, 12_345]])
        r = g.tick(src)
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch2, enabled=en), orc.make_packets(src, 0.0, ch2[src]), threads=8)
        m = r.merged()
        assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
        assert np.array_equal(m["outcome"], ref["pairs"]["outcome"])
        # dense move of everybody
        x2, y2 = x + rng.uniform(-1, 1, 20_000), y + rng.uniform(-1, 1, 20_000)
        g.update_positions(x2, y2)
        r = g.tick(src)
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x2, y2, z, channel=ch2, enabled=en), orc.make_packets(src, 0.0, ch2[src]), threads=8)
        m = r.merged()
        assert np.array_equal(m["rx"], ref["pairs"]["rx"]) and np.array_equal(m["outcome"], ref["pairs"]["outcome"])
        # a node leaves: the partition moves (Simulator.removeNode compacts the array, Simulator.java:279-310)
        keep = np.delete(np.arange(20_000), 9_999)
        src2 = pick_sources(74, 19_999, 200)
        g.set_nodes(x2[keep], y2[keep], channel=ch2[keep], enabled=en[keep])
        r = g.tick(src2)
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x2[keep], y2[keep], channel=ch2[keep], enabled=en[keep]),
                       orc.make_packets(src2, 0.0, ch2[keep][src2]), threads=8)
        m = r.merged()
        assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
        assert np.array_equal(m["outcome"], ref["pairs"]["outcome"])


def test_group_reference_media_and_errors(rsb, orc):
    from radio_sim_b200 import _abi as A
    x, y, z, ch = random_deployment(75, 3000, 400.0, channels=2)
    src = pick_sources(75, 3000, 40)
    with rsb.Group("udgm", devices=(0, 0), max_tx_per_tick=64) as g:
        g.set_nodes(x, y, z, channel=ch)
        m = g.tick(src).merged()
    ref = orc.tick(orc.params(orc.MODEL_UDGM), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src]))
    assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
    # stochastic reception: one shard draws by itself, several shards follow the one global draw order
    rxp = np.full(3000, 0.5)
    u = orc.jrandom_doubles(5, 20000)
    ref = orc.tick(orc.params(orc.MODEL_UDGM), orc.NodeSet(x, y, z, channel=ch, rxprob=rxp), orc.make_packets(src, 0.0, ch[src]), uniforms=u)
    for devices in ((0,), (0, 0), (0, 0, 0)):
        with rsb.Group("udgm", devices=devices, max_tx_per_tick=64) as g:
            g.set_nodes(x, y, z, channel=ch, rxprob=rxp)
            with pytest.raises(rsb.MediumError) as e:
                g.tick(src)                                # no uniforms: told how many the tick can consume
            assert e.value.code == A.ERR_UNIFORMS
            r = g.tick(src, uniforms=u)
        m = r.merged()
        assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
        assert np.array_equal(m["outcome"], ref["pairs"]["outcome"]) and r.n_uniforms_used == ref["n_uniforms_used"] > 100
    # more packets from one shard than the exchange buffers were sized for
    with rsb.Group("sinr", devices=(0, 0), max_tx_per_tick=8) as g:
        g.set_nodes(x, y, z, channel=ch)
        g.tick(src[:8])
        with pytest.raises(rsb.MediumError) as e:
            g.tick(np.arange(40, dtype=np.int32))
        assert e.value.code == A.ERR_CAPACITY


def test_group_with_fewer_nodes_than_devices(rsb, orc):
    """Three nodes on four handles: one shard owns nobody, another a single node; the lists still merge to the oracle's."""
    x, y = np.array([0.0, 20.0, 45.0]), np.zeros(3)
    for model, omodel in (("udgm", 1), ("sinr", 4)):
        with rsb.Group(model, devices=(0, 0, 0, 0), max_tx_per_tick=8, compact=True) as g:
            g.set_nodes(x, y)
            m = g.tick(np.array([0, 2], dtype=np.int32)).merged()
            rssi, recv = g.node_info_f64()
        ref = orc.tick(orc.params(omodel), orc.NodeSet(x, y), orc.make_packets(np.array([0, 2]), 0.0, 26))
        assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
        assert np.array_equal(m["outcome"], ref["pairs"]["outcome"]) and len(rssi) == 3


def test_group_udgm_success_ratio_and_tx_loss_on_several_shards(rsb, orc):
    """The reference's own stochastic medium sharded: UDGM with SUCCESS_RATIO_RX < 1 (which the reference applies to the
    transmission, UDGMRadioMedium.java:63-65) and per-node rx/tx probabilities, three handles, two ticks out of one
    java.util.Random stream."""
    from radio_sim_b200.javarandom import JavaRandom
    n, t = 3000, 90
    x, y, z, ch = random_deployment(88, n, 420.0, channels=2)
    rng = np.random.default_rng(88)
    rxprob = np.where(rng.uniform(size=n) < 0.4, 0.7, 1.0); rxprob[::41] = 0.0
    txprob = np.where(rng.uniform(size=n) < 0.4, 0.85, 1.0); txprob[::53] = 0.0
    nodes = orc.NodeSet(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob)
    p = orc.params(orc.MODEL_UDGM, success_ratio_rx=0.9)
    rnd, rnd_ref = JavaRandom(7), JavaRandom(7)
    with rsb.Group("udgm", devices=(0, 0, 0), max_tx_per_tick=128, success_ratio_rx=0.9) as g:
        g.set_nodes(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob)
        for tick in range(2):
            src = pick_sources(88 + tick, n, t)
            probe = g.tick(src, allow_short_uniforms=True)
            r = g.tick(src, uniforms=rnd.peek_doubles(probe.n_uniforms_used))
            rnd.skip_doubles(r.n_uniforms_used)
            ref = orc.tick(p, nodes, orc.make_packets(src, 0.0, ch[src]), uniforms=rnd_ref.peek_doubles(100000))
            rnd_ref.skip_doubles(ref["n_uniforms_used"])
            m = r.merged()
            assert r.status == 0 and r.n_uniforms_used == ref["n_uniforms_used"] > 50
            assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
            assert np.array_equal(m["outcome"], ref["pairs"]["outcome"])
            assert np.array_equal(r.tx_failed(), ref["tx_failed"]) and ref["tx_failed"].sum() > 0
        assert rnd.seed == rnd_ref.seed


@pytest.mark.parametrize("compact", [False, True])
def test_group_stochastic_reception_on_several_shards(rsb, orc, compact):
    """rx-loss / tx-loss < 1 with the receivers of ONE group on three handles: rmb_group_tick gathers the shards' draw
    counts itself and every shard walks the one java.util.Random stream -- same lists, same tx failures and the same
    number of uniforms consumed as the oracle's unsharded tick, through the adapter's two calls (the first, without
    uniforms, tells how many the tick can consume at most)."""
    from radio_sim_b200 import _abi as A
    from radio_sim_b200.javarandom import JavaRandom
    n, t = 2400, 140
    x, y, z, ch = random_deployment(61, n, 650.0)
    rng = np.random.default_rng(61)
    rxprob = np.where(rng.uniform(size=n) < 0.5, 0.9, 1.0); rxprob[::50] = 0.0
    txprob = np.where(rng.uniform(size=n) < 0.3, 0.8, 1.0); txprob[::97] = 0.0
    rxprob[:n // 3] = 1.0; txprob[:n // 3] = 1.0        # shard 0 alone is deterministic: it has to follow the others
    nodes = orc.NodeSet(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob)
    rnd, rnd_ref = JavaRandom(42), JavaRandom(42)
    with rsb.Group("sinr", devices=(0, 0, 0), max_tx_per_tick=256, compact=compact, f64_values=not compact) as g:
        g.set_nodes(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob)
        for tick in range(3):
            src = pick_sources(61 + tick, n, t) if tick < 2 else np.zeros(0, dtype=np.int32)
            probe = g.tick(src, allow_short_uniforms=True)
            if len(src):
                assert probe.status == A.ERR_UNIFORMS and probe.n_uniforms_used > 50
            r = g.tick(src, uniforms=rnd.peek_doubles(probe.n_uniforms_used))
            rnd.skip_doubles(r.n_uniforms_used)
            u = rnd_ref.peek_doubles(100000)
            ref = orc.tick(orc.params(orc.MODEL_SINR), nodes, orc.make_packets(src, 0.0, 26), uniforms=u)
            rnd_ref.skip_doubles(ref["n_uniforms_used"])
            assert r.status == 0 and r.n_uniforms_used == ref["n_uniforms_used"] <= probe.n_uniforms_used
            m = r.merged()
            assert np.array_equal(m["tx"], ref["pairs"]["tx"]) and np.array_equal(m["rx"], ref["pairs"]["rx"])
            assert np.array_equal(m["outcome"], ref["pairs"]["outcome"])
            assert np.array_equal(r.tx_failed(), ref["tx_failed"])
            if len(src):
                assert ref["tx_failed"].sum() > 0 and (ref["pairs"]["outcome"] == 1).sum() > 0
        # back to deterministic nodes: the group leaves the draws path (and consumes nothing)
        g.set_nodes(x, y, z, channel=ch)
        r = g.tick(pick_sources(61, n, t))
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(pick_sources(61, n, t), 0.0, 26))
        assert r.status == 0 and r.n_uniforms_used == 0 and np.array_equal(r.merged()["rx"], ref["pairs"]["rx"])
