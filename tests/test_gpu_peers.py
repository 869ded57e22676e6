"""GPU: the tick's tx records exchanged by the library itself over peer memory (rmb_peer_*),
instead of a collective run by the caller.  Two receiver shards live as two handles in ONE process on
cuda:0 (the same-process form of the exchange, as a single Java host driving several GPUs would use
it); every tick both shards push, then both tick, and the merged lists must equal the oracle's
unsharded tick.  The one-process-per-GPU form (CUDA IPC) is exercised by bench.py --gpus N."""
import numpy as np
import pytest

from helpers import random_deployment, pick_sources, SINR_TOL_DB

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rsb():
    import radio_sim_b200
    radio_sim_b200.load()
    return radio_sim_b200


def _shards(rsb, x, y, z, ch, world, max_slots):
    from radio_sim_b200.sharding import shard_bounds
    bounds = shard_bounds(len(x), world)
    meds = []
    for r, (lo, hi) in enumerate(bounds):
        m = rsb.Medium("sinr", f64_values=True, rx_base=lo, n_shards=world)
        m.set_nodes(x[lo:hi], y[lo:hi], z[lo:hi], channel=ch[lo:hi])
        meds.append(m)
    blobs = [m.peer_export(world, r, max_slots) for r, m in enumerate(meds)]
    for m in meds:
        m.peer_open(blobs)
    return bounds, meds


@pytest.mark.parametrize("world", [2, 3])
def test_peer_exchange_matches_the_unsharded_oracle(rsb, orc, world):
    from radio_sim_b200.sharding import TickPlan, merge_results
    n = 3000
    x, y, z, ch = random_deployment(41, n, 500.0, channels=2)
    bounds, meds = _shards(rsb, x, y, z, ch, world, max_slots=128)
    try:
        for tick in range(4):                          # both buffer parities, growing epochs, changing slot counts
            t = [90, 40, 0, 120][tick]
            src = pick_sources(41 + tick, n, t)
            power = np.random.default_rng(tick).choice([-7.0, 0.0], size=t)
            plan = TickPlan(src, bounds)
            for r, m in enumerate(meds):
                m.peer_push(plan.local_src[r], plan.local_fields(r, power, 0.0))
            parts = []
            for m in meds:
                res = m.tick_peers()
                assert res.n_tx == world * plan.tpad
                parts.append(dict(tx=res.tx, rx=res.rx, outcome=res.outcome, rssi=res.rssi_dbm_f64, sinr=res.sinr_db_f64))
            merged = merge_results(plan, parts)
            full = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, power, ch[src]))["pairs"]
            assert np.array_equal(merged["tx"], full["tx"]) and np.array_equal(merged["rx"], full["rx"])
            assert np.array_equal(merged["outcome"], full["outcome"])
            assert np.array_equal(merged["rssi"], full["rssi_dbm"])
            if len(full["tx"]):
                fin = np.isfinite(full["sinr_db"])
                assert np.abs(merged["sinr"][fin] - full["sinr_db"][fin]).max() < SINR_TOL_DB
    finally:
        for m in meds:
            m.close()


def test_peer_call_order_is_checked(rsb):
    x, y, z, ch = random_deployment(5, 200, 100.0)
    with rsb.Medium("sinr") as m:
        m.set_nodes(x, y)
        with pytest.raises(rsb.MediumError):           # push before export/open
            m.peer_push(np.array([1], dtype=np.int32))
        blob = m.peer_export(1, 0, 8)
        m.peer_open([blob])
        with pytest.raises(rsb.MediumError):           # tick without a push
            m.tick_peers()
        with pytest.raises(rsb.MediumError):           # more slots than exported
            m.peer_push(np.arange(9, dtype=np.int32))
        m.peer_push(np.array([3, 7, -1], dtype=np.int32))
        with pytest.raises(rsb.MediumError):           # second push without a tick
            m.peer_push(np.array([3], dtype=np.int32))
        r = m.tick_peers()
        assert r.n_tx == 3 and r.n_pairs > 0
        ref = m.tick(np.array([3, 7], dtype=np.int32))
        assert np.array_equal(r.rx, ref.rx) and np.array_equal(r.outcome, ref.outcome)


def test_fused_exchange_in_the_preparation_kernel(rsb, orc):
    """rmb_peer_exchange_tick_enqueue: the preparation kernel itself pushes the records, raises the flags and
    waits.  Two cooperative kernels of one process do not run side by side on ONE device, so on this single-GPU
    box the fused form is exercised with a world of one rank (push -> own gather buffer -> flag -> wait); with
    one GPU per rank it is what bench.py --gpus N runs, checked there against the single-GPU pair count."""
    import torch
    n = 3000
    x, y, z, ch = random_deployment(43, n, 500.0)
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        m.peer_open([m.peer_export(1, 0, 128)])
        for tick in range(3):
            src = pick_sources(50 + tick, n, 70)
            padded = np.concatenate([src[:40], [-1, -1], src[40:]]).astype(np.int32)      # padding keeps its arrival slot
            rec = torch.zeros(padded.size * 48, dtype=torch.uint8, device="cuda")
            m.build_tx_records(padded, rec.data_ptr())
            m.peer_exchange_tick_enqueue(padded.size, rec.data_ptr())
            counts = m.tick_finish()
            res = m.fetch()
            full = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, 26))["pairs"]
            assert counts.n_pairs == len(full["tx"])
            slot_of = np.flatnonzero(padded >= 0)
            assert np.array_equal(res.tx, slot_of[full["tx"]]) and np.array_equal(res.rx, full["rx"])
            assert np.array_equal(res.outcome, full["outcome"]) and np.array_equal(res.rssi_dbm_f64, full["rssi_dbm"])


@pytest.mark.parametrize("world", [1, 3])
def test_stochastic_reception_on_sharded_receivers_follows_the_global_draw_order(rsb, orc, world):
    """rx-loss / tx-loss < 1 with the receivers sharded: rmb_tick_draws_begin -> gather of the per-packet draw
    counts -> rmb_tick_draws_end; outcomes, tx failures and the number of uniforms consumed must equal the
    oracle's unsharded tick fed with the same java.util.Random(42) stream."""
    import torch
    from radio_sim_b200.sharding import shard_bounds, TickPlan, merge_results
    n, t = 2400, 140
    x, y, z, ch = random_deployment(61, n, 650.0)
    rng = np.random.default_rng(61)
    rxprob = np.where(rng.uniform(size=n) < 0.5, 0.9, 1.0); rxprob[::50] = 0.0
    txprob = np.where(rng.uniform(size=n) < 0.3, 0.8, 1.0); txprob[::97] = 0.0
    if world > 1:
        rxprob[:n // world] = 1.0                  # shard 0 alone would see no probability < 1: it still has to follow
    u = orc.jrandom_doubles(42, 100000)
    bounds = shard_bounds(n, world)
    meds = []
    for r, (lo, hi) in enumerate(bounds):
        m = rsb.Medium("sinr", f64_values=True, rx_base=lo, n_shards=world)
        m.set_nodes(x[lo:hi], y[lo:hi], z[lo:hi], channel=ch[lo:hi], rxprob=rxprob[lo:hi], txprob=txprob[lo:hi])
        meds.append(m)
    blobs = [m.peer_export(world, r, 256) for r, m in enumerate(meds)]
    for m in meds:
        m.peer_open(blobs)
    try:
        for tick in range(2):
            src = pick_sources(61 + tick, n, t)
            plan = TickPlan(src, bounds)
            n_rec = world * plan.tpad
            for r, m in enumerate(meds):
                m.peer_push(plan.local_src[r])
            counts = torch.zeros(world * n_rec, dtype=torch.int32, device="cuda")      # [world][n_rec], "all-gathered" in place
            for r, m in enumerate(meds):
                m.tick_peers_draws_begin(counts.data_ptr() + 4 * r * n_rec)
            order = torch.from_numpy(plan.order).cuda()
            torch.cuda.synchronize()
            parts, used, failed = [], [], []
            for r, m in enumerate(meds):
                res = m.tick_draws_end(counts.data_ptr(), world, r, u, d_order=order.data_ptr(), n_order=t)
                parts.append(dict(tx=res.tx, rx=res.rx, outcome=res.outcome, rssi=res.rssi_dbm_f64))
                used.append(res.n_uniforms_used); failed.append(res.tx_failed.copy())
            merged = merge_results(plan, parts)
            ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob),
                           orc.make_packets(src, 0.0, 26), uniforms=u)
            full = ref["pairs"]
            assert ref["n_uniforms_used"] > 50 and ref["tx_failed"].sum() > 0
            assert all(v == ref["n_uniforms_used"] for v in used)
            valid = np.flatnonzero(plan.packet_of >= 0)
            for f in failed:                            # every shard walks the same tx draws
                got = np.zeros(t, dtype=np.uint8); got[plan.packet_of[valid]] = f[valid]
                assert np.array_equal(got, ref["tx_failed"])
            assert np.array_equal(merged["tx"], full["tx"]) and np.array_equal(merged["rx"], full["rx"])
            assert np.array_equal(merged["outcome"], full["outcome"]) and np.array_equal(merged["rssi"], full["rssi_dbm"])
    finally:
        for m in meds:
            m.close()


@pytest.mark.parametrize("model", ["sinr", "udgm"])
def test_peer_tick_one_call_per_rank_from_concurrent_threads(rsb, orc, model):
    """rmb_peer_tick: push + tick in one call, every rank inside it at the same time (one host thread per handle, as
    rmb_group_* runs them).  Handles that share a device take the separate push inside the call; mixed with the
    two-call form in between, pushes and ticks stay in step."""
    import threading
    from radio_sim_b200.sharding import shard_bounds, TickPlan, merge_results
    n, world = 3000, 3
    x, y, z, ch = random_deployment(43, n, 500.0, channels=2)
    bounds = shard_bounds(n, world)
    meds = []
    for r, (lo, hi) in enumerate(bounds):
        m = rsb.Medium(model, f64_values=True, rx_base=lo, n_shards=world)
        m.set_nodes(x[lo:hi], y[lo:hi], z[lo:hi], channel=ch[lo:hi])
        meds.append(m)
    blobs = [m.peer_export(world, r, 128) for r, m in enumerate(meds)]
    for m in meds:
        m.peer_open(blobs)
    omodel = orc.MODEL_SINR if model == "sinr" else orc.MODEL_UDGM
    try:
        for tick in range(5):
            t = [90, 40, 0, 120, 7][tick]
            src = pick_sources(43 + tick, n, t)
            plan = TickPlan(src, bounds)
            parts, errors = [None] * world, []
            if tick == 3:                                   # the two-call form between one-call ticks
                for r, m in enumerate(meds):
                    m.peer_push(plan.local_src[r])
                for r, m in enumerate(meds):
                    parts[r] = m.tick_peers()
            else:
                def run(r):
                    try:
                        parts[r] = meds[r].peer_tick(plan.local_src[r])
                    except Exception as e:                  # noqa: BLE001 - reported below
                        errors.append((r, e))
                threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
                for th in threads:
                    th.start()
                for th in threads:
                    th.join()
                assert not errors, errors
            merged = merge_results(plan, [dict(tx=p.tx, rx=p.rx, outcome=p.outcome, rssi=p.rssi_dbm_f64) for p in parts])
            full = orc.tick(orc.params(omodel), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, ch[src]))["pairs"]
            assert np.array_equal(merged["tx"], full["tx"]) and np.array_equal(merged["rx"], full["rx"])
            assert np.array_equal(merged["outcome"], full["outcome"]) and np.array_equal(merged["rssi"], full["rssi_dbm"])
    finally:
        for m in meds:
            m.close()


def test_peer_tick_fused_form_with_a_world_of_one(rsb, orc):
    """rmb_peer_tick takes the fused form (records pushed by the tick's first kernel) only when no peer shares the
    handle's device, so on a one-GPU box it is exercised with a world of one rank, like the enqueued form above; with
    one GPU per rank it is what bench.py --gpus N times and verifies (digest of the merged lists)."""
    n = 3000
    x, y, z, ch = random_deployment(44, n, 500.0)
    with rsb.Medium("sinr", f64_values=True) as m:
        m.set_nodes(x, y, z, channel=ch)
        m.peer_open([m.peer_export(1, 0, 128)])
        for tick in range(4):
            src = pick_sources(60 + tick, n, [70, 5, 0, 110][tick])
            padded = np.concatenate([src[:3], [-1, -1], src[3:]]).astype(np.int32)
            res = m.peer_tick(padded)
            full = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch), orc.make_packets(src, 0.0, 26))["pairs"]
            slot_of = np.flatnonzero(padded >= 0)
            assert res.n_pairs == len(full["tx"])
            assert np.array_equal(res.tx, slot_of[full["tx"]]) and np.array_equal(res.rx, full["rx"])
            assert np.array_equal(res.outcome, full["outcome"]) and np.array_equal(res.rssi_dbm_f64, full["rssi_dbm"])
