"""-m gpu: the SINR medium (K1 sweep + exact resolve) against the CPU oracle, through the C-ABI."""
import numpy as np
import pytest
from helpers import random_deployment, pick_sources, compare, SINR_TOL_DB

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rsb():
    import radio_sim_b200
    return radio_sim_b200


def run_case(rsb, orc, x, y, z, ch, src, power=None, pkt_channel=None, enabled=None, rxprob=None, txprob=None,
             uniforms=None, variant=0, slots=0, **params):
    with rsb.Medium("sinr", f64_values=True, **params) as m:
        if variant or slots:
            m.set_tuning(variant, slots)
        m.set_nodes(x, y, z, channel=ch, enabled=enabled, rxprob=rxprob, txprob=txprob)
        r = m.tick(src, power_dbm=power, channel=pkt_channel, uniforms=uniforms)
        rssi, recv = m.node_info()
    nodes = orc.NodeSet(x, y, z, channel=ch, enabled=enabled, rxprob=rxprob, txprob=txprob)
    pw = np.zeros(len(src)) if power is None else power
    pc = ch[src] if pkt_channel is None else pkt_channel
    ref = orc.tick(orc.params(orc.MODEL_SINR, **params), nodes, orc.make_packets(src, pw, pc), uniforms=uniforms,
                   want_total=True, threads=8)
    compare(r, ref)
    if uniforms is not None:
        assert r.n_uniforms_used == ref["n_uniforms_used"]
    assert np.abs(rssi.astype(np.float64) - ref["total_dbm"]).max() < 1e-3
    heard = np.zeros(len(x), dtype=np.uint8)
    heard[ref["pairs"]["rx"]] = 1
    assert np.array_equal(recv, heard)
    return r, ref


# sweep variants, then tuning bits: 0x10000 stand-alone resolve kernel, 0x20000 / 0x80000 final-wave split multiplier 1 / 4,
# 0x100000 separate recheck / ordering / emit kernels instead of the one finish kernel
@pytest.mark.parametrize("variant", [0, 2, 3, 4, 5, 6, 7, 8, 0x10000, 0x20000, 0x80000, 0x100000, 0x110000])
def test_config2_1k_nodes_50_tx(rsb, orc, variant):
    x, y, z, ch = random_deployment(1, 1000, 1000.0)
    r, _ = run_case(rsb, orc, x, y, z, ch, pick_sources(1, 1000, 50), variant=variant)
    assert r.n_pairs > 0 and r.n_delivered > 0 and r.n_interfered > 0


def test_dense_all_hear_all_grows_slots(rsb, orc):
    x, y, z, ch = random_deployment(2, 600, 60.0)
    r, _ = run_case(rsb, orc, x, y, z, ch, pick_sources(2, 600, 40))
    assert r.near_slots >= 32


@pytest.mark.parametrize("n,tuning", [(50, 0), (300, 0), (900, 0), (1500, 0), (300, 0x100000)])
def test_every_bucket_ranking_path(rsb, orc, n, tuning):
    """All nodes within hearing distance: a packet's bucket holds ~n pairs -- one warp (<= 64), counting in a block
    (<= 512), bitonic sort in a block (<= 2048) in finish.cu."""
    x, y, z, ch = random_deployment(20 + n, n, 40.0)
    r, _ = run_case(rsb, orc, x, y, z, ch, pick_sources(n, n, 12), variant=tuning)
    assert r.n_pairs >= 12 * (n - 12)


def test_three_dimensional(rsb, orc):
    x, y, z, ch = random_deployment(3, 3000, 500.0, z_spread=40.0)
    run_case(rsb, orc, x, y, z, ch, pick_sources(3, 3000, 100))


def test_16_channels_and_packet_channel_override(rsb, orc):
    x, y, z, ch = random_deployment(4, 5000, 800.0, channels=16)
    src = pick_sources(4, 5000, 300)
    pc = ch[src].copy()
    pc[::7] = 11            # per-packet "wireless-channel" override (SimulatorJSONHandler.java:87-90)
    pc[5] = 99              # nobody listens there
    run_case(rsb, orc, x, y, z, ch, src, pkt_channel=pc)


def test_mixed_power_disabled_and_duplicates(rsb, orc):
    x, y, z, ch = random_deployment(5, 4000, 700.0)
    src = pick_sources(5, 4000, 200)
    src[10] = src[3]        # one node sends two frames in the tick
    power = np.random.default_rng(5).choice([-25.0, -15.0, -7.0, 0.0, 5.0], size=200)
    enabled = (np.random.default_rng(6).uniform(size=4000) > 0.1).astype(np.uint8)
    x[17], y[17] = x[src[0]], y[src[0]]      # co-located with a transmitter: d = 0 -> clamped to d0
    run_case(rsb, orc, x, y, z, ch, src, power=power, enabled=enabled)


@pytest.mark.parametrize("n_exp", [2.0, 3.0, 4.0, 2.7])
def test_path_loss_exponents(rsb, orc, n_exp):
    x, y, z, ch = random_deployment(7, 3000, 1500.0)
    run_case(rsb, orc, x, y, z, ch, pick_sources(7, 3000, 64), pl_exponent=n_exp, sensitivity_dbm=-100.0 if n_exp < 3 else -95.0)


def test_stochastic_reception_uses_host_uniforms(rsb, orc):
    x, y, z, ch = random_deployment(8, 2000, 600.0)
    rng = np.random.default_rng(8)
    rxprob = np.where(rng.uniform(size=2000) < 0.5, 0.9, 1.0)
    rxprob[::50] = 0.0
    txprob = np.where(rng.uniform(size=2000) < 0.3, 0.8, 1.0)
    txprob[::97] = 0.0
    u = orc.jrandom_doubles(42, 100000)
    r, ref = run_case(rsb, orc, x, y, z, ch, pick_sources(8, 2000, 150), rxprob=rxprob, txprob=txprob, uniforms=u)
    assert ref["n_uniforms_used"] > 100 and ref["tx_failed"].sum() > 0


def test_empty_and_tiny(rsb, orc):
    x, y, z, ch = random_deployment(9, 3, 10.0)
    r, _ = run_case(rsb, orc, x, y, z, ch, np.zeros(0, dtype=np.int32))
    assert r.n_pairs == 0
    run_case(rsb, orc, x, y, z, ch, np.array([0, 1, 2], dtype=np.int32))


def test_no_enabled_receiver_at_all(rsb, orc):
    """Every radio disabled: no sweep, no resolve, hence no segment offsets for the tick -- the tail of the tick must
    not walk stale ones.  Enabling the radios afterwards gives the ordinary tick."""
    x, y, z, ch = random_deployment(14, 800, 150.0)
    src = pick_sources(14, 800, 20)
    off = np.zeros(800, dtype=np.uint8)
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        first = m.tick(src)                                   # leaves segment offsets of a busy tick behind
        assert first.n_pairs > 1000
        m.update_radio(enabled=off)
        r = m.tick(src)
        assert r.n_pairs == 0 and r.n_delivered == 0 and r.n_interfered == 0 and r.n_decisions == 0
        m.update_radio(enabled=np.ones(800, dtype=np.uint8))
        compare(m.tick(src), orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src])))
    run_case(rsb, orc, x, y, z, ch, src, enabled=off)         # and from a fresh handle, against the oracle


def test_guard_band_forces_exact_recheck(rsb, orc):
    """Place an interferer so that one receiver's SINR sits on the 4 dB threshold: the decision must
    come from the exact double path and agree with the oracle on both sides of the threshold."""
    base = np.array([[0.0, 0.0], [20.0, 0.0]])            # tx A at origin, receiver at 20 m
    # find the interferer distance that gives SINR == threshold for the receiver (bisection on the oracle)
    def sinr_at(d):
        xs = np.array([0.0, 20.0, 20.0 + d]); ys = np.zeros(3)
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(xs, ys), orc.make_packets([0, 2], 0.0, 26))
        p = ref["pairs"]
        return float(p["sinr_db"][(p["tx"] == 0) & (p["rx"] == 1)][0])
    lo, hi = 1.0, 500.0
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if sinr_at(mid) < 4.0:
            lo = mid
        else:
            hi = mid
    total_recheck = 0
    for d in [lo, hi, np.nextafter(lo, 0), np.nextafter(hi, 1e9), lo - 1e-9, hi + 1e-9]:
        xs = np.array([0.0, 20.0, 20.0 + d]); ys = np.zeros(3)
        r, _ = run_case(rsb, orc, xs, ys, np.zeros(3), np.full(3, 26, dtype=np.int32), np.array([0, 2], dtype=np.int32))
        total_recheck += r.n_recheck
    assert total_recheck >= 4


def test_threshold_among_many_interferers_takes_the_sequential_sum(rsb, orc):
    """400 co-channel packets, one receiver whose SINR is bisected onto the threshold to the last ulp of
    the interferer distance: the tree sum cannot decide (|SINR - 4 dB| ~ 1e-13 dB), so the strictly
    sequential arrival-order sum must run and reproduce the oracle's decision on both sides."""
    rng = np.random.default_rng(11)
    n, t = 3000, 400
    x, y = rng.uniform(0, 4000, n), rng.uniform(0, 4000, n)
    x[0], y[0] = 2000.0, 2000.0            # the wanted transmitter
    x[1], y[1] = 2025.0, 2000.0            # the receiver, 25 m away
    src = np.concatenate([[0, 2], 3 + rng.permutation(n - 3)[:t - 2]]).astype(np.int32)
    z, ch = np.zeros(n), np.full(n, 26, dtype=np.int32)
    tiny = orc.params(orc.MODEL_SINR)

    def decodable(d):
        x[2], y[2] = 2025.0 + d, 2000.0    # the dominant interferer
        ref = orc.tick(tiny, orc.NodeSet(x, y), orc.make_packets(src, 0.0, 26), threads=8)
        p = ref["pairs"]
        sel = (p["tx"] == 0) & (p["rx"] == 1)
        assert sel.sum() == 1
        return bool(p["outcome"][sel][0] == 2)
    lo, hi = 5.0, 400.0
    assert not decodable(lo) and decodable(hi)
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if mid == lo or mid == hi:
            break
        if decodable(mid):
            hi = mid
        else:
            lo = mid
    seq = 0
    for d in [lo, hi]:
        x[2], y[2] = 2025.0 + d, 2000.0
        with rsb.Medium("sinr", f64_values=True) as m:
            m.set_nodes(x, y, z, channel=ch)
            r = m.tick(src)
            seq += m.timing()["recheck_sequential"]
        ref = orc.tick(tiny, orc.NodeSet(x, y), orc.make_packets(src, 0.0, 26), threads=8)
        compare(r, ref)
    assert seq >= 1


def test_packet_heard_by_thousands_takes_the_radix_fallback(rsb, orc):
    """3 packets heard by ~6000 receivers each: buckets larger than the in-block sort (2048) must be ordered
    by the radix-sort fallback; record capacity (2 x receivers at first) has to grow as well."""
    x, y, z, ch = random_deployment(12, 6000, 45.0)
    r, _ = run_case(rsb, orc, x, y, z, ch, np.array([5, 17, 4000], dtype=np.int32))
    assert r.n_pairs > 3 * 5000


def test_position_and_radio_updates(rsb, orc):
    x, y, z, ch = random_deployment(10, 3000, 600.0, channels=4)
    src = pick_sources(10, 3000, 100)
    rng = np.random.default_rng(10)
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        m.tick(src)
        # dense move
        x2, y2 = x + rng.uniform(-0.5, 0.5, 3000), y + rng.uniform(-0.5, 0.5, 3000)
        m.update_positions(x2, y2)
        compare(m.tick(src), orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x2, y2, None, channel=ch), orc.make_packets(src, 0.0, ch[src])))
        # sparse move with z, channel change, disable
        idx = np.array([5, 77, 1234], dtype=np.int32)
        x2[idx] += 3.0; z2 = np.zeros(3000); z2[idx] = 2.5
        m.update_positions(x2[idx], y2[idx], z2[idx], idx=idx)
        ch2 = ch.copy(); ch2[100:200] = 26
        en = np.ones(3000, dtype=np.uint8); en[300:400] = 0
        m.update_radio(idx=np.arange(100, 200, dtype=np.int32), channel=ch2[100:200])
        m.update_radio(idx=np.arange(300, 400, dtype=np.int32), enabled=en[300:400])
        pw = np.full(3000, -3.0)
        m.update_radio(txpower_dbm=pw)
        r = m.tick(src)
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x2, y2, z2, channel=ch2, enabled=en), orc.make_packets(src, -3.0, ch2[src]))
        compare(r, ref)


def test_mobility_over_many_dense_updates(rsb, orc):
    """BASELINE configs[4] in small: every node moves every tick, fresh sources per tick.  The library re-sorts its
    work order only on every 8th dense update; the steps here are large (nodes cross the whole area), so a stale
    order is as bad as it gets -- results must not depend on it."""
    n = 4000
    x, y, z, ch = random_deployment(12, n, 700.0, channels=16)
    rng = np.random.default_rng(12)
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        for tick in range(11):
            step = 0.5 if tick % 3 else 300.0
            x = np.clip(x + rng.uniform(-step, step, n), 0.0, 700.0)
            y = np.clip(y + rng.uniform(-step, step, n), 0.0, 700.0)
            m.update_positions(x, y)
            src = pick_sources(100 + tick, n, 160)
            ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, None, channel=ch), orc.make_packets(src, 0.0, ch[src]))
            compare(m.tick(src), ref)


def test_enqueued_tick_equals_the_synchronous_one(rsb, orc):
    """rmb_tick_enqueue + rmb_tick_finish + rmb_fetch: same lists as rmb_tick; an enqueued tick whose
    near-slot capacity is exceeded is redone when it is settled."""
    import torch
    x, y, z, ch = random_deployment(77, 4000, 600.0)
    src = pick_sources(77, 4000, 120)
    ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, 26))
    rec = torch.zeros(120 * 48, dtype=torch.uint8, device="cuda")
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        m.set_tuning(0, 1)                             # one near slot: the first enqueued tick overflows
        m.build_tx_records(src, rec.data_ptr())
        for _ in range(2):
            m.tick_enqueue(120, rec.data_ptr())
            counts = m.tick_finish()
            assert counts.n_pairs == ref["n_pairs"] and counts.n_decisions == ref["n_decisions"]
            compare(m.fetch(), ref)
        m.tick_enqueue(120, rec.data_ptr())            # settled implicitly by the next call
        compare(m.fetch(), ref)


@pytest.mark.parametrize("model", ["sinr", "udgm"])
def test_compact_result_form(rsb, orc, model):
    """RMB_FLAG_COMPACT: rx / rssi / outcome per row + tx_first per packet, exactly what the reference's
    generateReceptionEvents calls need; the tx and sinr columns are not transferred."""
    x, y, z, ch = random_deployment(5, 3000, 400.0)
    src = pick_sources(5, 3000, 80)
    with rsb.Medium(model) as full, rsb.Medium(model, compact=True) as cm:
        full.set_nodes(x, y, z, channel=ch)
        cm.set_nodes(x, y, z, channel=ch)
        a, b = full.tick(src), cm.tick(src)
    assert b.tx.size == 0 and b.sinr_db.size == 0 and b.tx_first is not None
    assert b.tx_first[0] == 0 and b.tx_first[-1] == a.n_pairs == b.n_pairs
    assert np.array_equal(b.tx_column(), a.tx) and np.array_equal(b.rx, a.rx)
    assert np.array_equal(b.outcome, a.outcome) and np.array_equal(b.rssi_dbm, a.rssi_dbm)
    assert np.array_equal(b.tx_failed, a.tx_failed)


def test_half_duplex_flag_against_the_oracle(rsb, orc):
    """rmb_config.half_duplex (SURVEY 8(a) open row): a node that is the source of any packet of the tick passes no
    receiver filter, reports the noise floor, and the decision count drops accordingly; default = the reference."""
    x, y, z, ch = random_deployment(61, 4000, 500.0, channels=2)
    src = pick_sources(61, 4000, 120)
    src[5] = src[4]                                        # one node with two packets in the tick
    for hd in (0, 1):
        with rsb.Medium("sinr", f64_values=True, half_duplex=hd) as m:
            m.set_nodes(x, y, z, channel=ch)
            r = m.tick(src)
            r2 = m.tick(src[:60])                          # the stamps of the previous tick must not leak
            rssi, _ = m.node_info_f64()
        p = orc.params(orc.MODEL_SINR, half_duplex=hd)
        ref = orc.tick(p, orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src]))
        ref2 = orc.tick(p, orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src[:60], 0.0, ch[src[:60]]), want_total=True)
        compare(r, ref)
        compare(r2, ref2)
        assert np.abs(rssi - ref2["total_dbm"]).max() < SINR_TOL_DB
        if hd:
            assert not np.isin(r.rx, src).any()
            assert np.all(rssi[src[:60]] == -100.0) and rssi[src[100]] > -100.0
    for model in ("udgm", "null", "udgm-constant"):
        with rsb.Medium(model, f64_values=True, half_duplex=1) as m:
            m.set_nodes(x, y, z, channel=ch)
            r = m.tick(src)
        ref = orc.tick(orc.params({"udgm": 1, "null": 0, "udgm-constant": 2}[model], half_duplex=1),
                       orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src]))
        compare(r, ref)


def test_limits_are_errors_not_wrong_results(rsb):
    """max_pairs beyond int32, N2N on sharded receivers, and a foreign record with 0 < txprob < 1 on a
    shard whose own nodes are deterministic."""
    import ctypes as C
    from radio_sim_b200 import _abi as A
    with pytest.raises(rsb.MediumError) as e:
        rsb.Medium("sinr", max_pairs=1 << 31)
    assert e.value.code == A.ERR_INVALID
    with pytest.raises(rsb.MediumError) as e:
        rsb.Medium("n2n", n_shards=2)
    assert e.value.code == A.ERR_UNSUPPORTED
    import torch
    x, y, z, ch = random_deployment(62, 2000, 300.0)
    for model in ("sinr", "udgm"):
        with rsb.Medium(model, rx_base=0, n_shards=2) as m:
            m.set_nodes(x[:1000], y[:1000])
            recs = torch.zeros(4 * 48, dtype=torch.uint8, device="cuda")
            m.build_tx_records(np.array([3, 10, 20, 30], dtype=np.int32), recs.data_ptr())
            r = np.frombuffer(recs.cpu().numpy().tobytes(), dtype=[("x", "f8"), ("y", "f8"), ("z", "f8"), ("p", "f8"), ("txprob", "f8"),
                                                             ("ch", "i4"), ("src", "i4")]).copy()
            r["src"][3] = 1500; r["x"][3] = x[1500]; r["y"][3] = y[1500]; r["txprob"][3] = 0.5     # as another shard built it
            recs.copy_(torch.from_numpy(np.frombuffer(r.tobytes(), dtype=np.uint8).copy()))
            with pytest.raises(rsb.MediumError) as e:
                m.tick_records(4, recs.data_ptr())
            assert e.value.code == A.ERR_UNSUPPORTED
            r["txprob"][3] = 1.0
            recs.copy_(torch.from_numpy(np.frombuffer(r.tobytes(), dtype=np.uint8).copy()))
            assert m.tick_records(4, recs.data_ptr()).n_tx == 4


def test_positions_uploaded_beside_the_running_tick(rsb, orc):
    """rmb_update_positions_async: tick k is enqueued, the positions of tick k+1 are uploaded while it runs and swapped
    in when the handle is next settled; every tick must see exactly its own positions (per-tick mobility, C5)."""
    import torch
    x, y, z, ch = random_deployment(81, 30_000, 1500.0, channels=4)
    rng = np.random.default_rng(82)
    n, t = len(x), 300
    px = [torch.empty(n, dtype=torch.float64).pin_memory().numpy() for _ in range(2)]
    py = [torch.empty(n, dtype=torch.float64).pin_memory().numpy() for _ in range(2)]
    px[0][:], py[0][:] = x, y
    rec = torch.zeros(t * 48, dtype=torch.uint8, device="cuda")
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        for k in range(12):
            cur, nxt = k % 2, (k + 1) % 2
            src = rng.permutation(n)[:t].astype(np.int32)
            m.build_tx_records(src, rec.data_ptr())                  # settles: positions of tick k are the node table now
            m.tick_enqueue(t, rec.data_ptr())
            px[nxt][:] = np.clip(px[cur] + rng.uniform(-0.5, 0.5, n), 0, 1500.0)
            py[nxt][:] = np.clip(py[cur] + rng.uniform(-0.5, 0.5, n), 0, 1500.0)
            m.update_positions_async(px[nxt], py[nxt])               # beside the running tick
            cnt = m.tick_finish()
            r = m.fetch()
            if k == 5:
                rssi, _ = m.node_info_f64()                          # of tick 5, although tick 6's positions are in already
            ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(px[cur].copy(), py[cur].copy(), z, channel=ch),
                           orc.make_packets(src, 0.0, ch[src]), threads=8, want_total=True)
            assert cnt.n_pairs == ref["n_pairs"]
            compare(r, ref)
            if k == 5:
                assert np.abs(rssi - ref["total_dbm"]).max() < SINR_TOL_DB
        # a staged set that nobody consumed before the node table changes is void
        m.update_positions_async(px[0], py[0])
        m.set_nodes(x[:1000], y[:1000], channel=ch[:1000])
        r = m.tick(np.arange(20, dtype=np.int32))
        ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x[:1000], y[:1000], channel=ch[:1000]),
                       orc.make_packets(np.arange(20), 0.0, ch[:20]))
        compare(r, ref)


@pytest.mark.parametrize("compact", [False, True])
def test_lists_copied_behind_the_tick_on_a_guessed_length(rsb, orc, compact):
    """The synchronous calls copy the result slab home behind the finish kernel on the previous tick's length (+ 1/8):
    ticks that grow, shrink, repeat and go empty must all come back whole (a short guess is topped up by a second copy)."""
    x, y, z, ch = random_deployment(91, 6000, 500.0)
    nodes = orc.NodeSet(x, y, z, channel=ch)
    p = orc.params(orc.MODEL_SINR)
    with rsb.Medium("sinr", f64_values=not compact, compact=compact) as m:
        m.set_nodes(x, y, z, channel=ch)
        for k, n_tx in enumerate([3, 3, 200, 200, 1, 0, 40, 45, 600, 2]):
            src = pick_sources(1000 + k, 6000, n_tx) if n_tx else np.zeros(0, dtype=np.int32)
            r = m.tick(src)
            ref = orc.tick(p, nodes, orc.make_packets(src, 0.0, ch[src]), threads=8)
            assert r.n_pairs == ref["n_pairs"]
            if compact:
                assert np.array_equal(r.tx_column(), ref["pairs"]["tx"]) and np.array_equal(r.rx, ref["pairs"]["rx"])
                assert np.array_equal(r.outcome, ref["pairs"]["outcome"])
                assert r.tx_first[-1] == ref["n_pairs"] and len(r.tx_first) == n_tx + 1
                assert np.array_equal(r.tx_failed, ref["tx_failed"])
            else:
                compare(r, ref)


def test_more_packets_than_the_finish_kernel_scans(rsb, orc):
    """Above 16 383 packets per tick the per-packet segment offsets are scanned in the sweep's epilogue instead of the
    finish kernel's shared memory (finish.cu finish_scans): 20 000 of 30 000 nodes transmit at once."""
    x, y, z, ch = random_deployment(123, 30000, 3000.0)
    ch[:] = 26
    src = pick_sources(123, 30000, 20000)
    r, ref = run_case(rsb, orc, x, y, z, ch, src)
    assert r.n_tx == 20000 and ref["n_pairs"] > 100000
    with rsb.Medium("sinr", compact=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        c = m.tick(src)
        assert np.array_equal(c.tx_column(), ref["pairs"]["tx"]) and np.array_equal(c.rx, ref["pairs"]["rx"])
        assert np.array_equal(c.outcome, ref["pairs"]["outcome"])
