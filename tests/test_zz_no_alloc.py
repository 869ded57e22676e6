"""-m gpu, collected last: rmb_udgm_tick_static, the entry point that allocates nothing (its scratch is a file-scope
__device__ array; no counterpart in the reference).  Called through the
C-ABI with the caller's device buffers and checked row for row against rmb_tick on a UDGM handle -- index lists, outcomes
and float RSSI bit-exact, as everywhere in tests/ -- on two calls in a row that share the scratch (the second tick is
smaller and different, so anything left over from the first would show)."""
import ctypes as C
import numpy as np
import pytest

from helpers import random_deployment, pick_sources

pytestmark = pytest.mark.gpu


def _static_tick(lib, torch, x, y, z, ch, enabled, src, power, pkt_channel, R, capacity, want_tx=True):
    """-> (rc, error text, dict of host arrays); every buffer is allocated HERE, by the caller"""
    dev = torch.device("cuda", 0)
    up = lambda a, dt: None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    ptr = lambda t: C.c_void_p(0 if t is None else t.data_ptr())
    d_x, d_y, d_z = up(x, np.float64), up(y, np.float64), up(z, np.float64)
    d_ch, d_en = up(ch, np.int32), up(enabled, np.uint8)
    d_src, d_pw, d_pc = up(src, np.int32), up(power, np.float64), up(pkt_channel, np.int32)
    t = len(src)
    cap = max(int(capacity), 1)
    o_tx = torch.full((cap,), -7, dtype=torch.int32, device=dev) if want_tx else None
    o_rx = torch.full((cap,), -7, dtype=torch.int32, device=dev)
    o_rssi = torch.zeros(cap, dtype=torch.float32, device=dev)
    o_out = torch.zeros(cap, dtype=torch.uint8, device=dev)
    o_first = torch.full((t + 1,), -7, dtype=torch.int32, device=dev)
    o_n = torch.full((1,), -7, dtype=torch.int64, device=dev)
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    rc = lib.rmb_udgm_tick_static(0, None, len(x), ptr(d_x), ptr(d_y), ptr(d_z), ptr(d_ch), ptr(d_en), t, ptr(d_src), ptr(d_pw),
                                  ptr(d_pc), float(R), int(capacity), ptr(o_tx), ptr(o_rx), ptr(o_rssi), ptr(o_out), ptr(o_first),
                                  ptr(o_n), err, 256)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(0)[0]
    out = None
    if rc == 0:
        m = int(o_n.item())
        k = min(m, int(capacity))
        out = dict(n_pairs=m, rx=o_rx[:k].cpu().numpy(), rssi=o_rssi[:k].cpu().numpy(), outcome=o_out[:k].cpu().numpy(),
                   tx=o_tx[:k].cpu().numpy() if want_tx else None, tx_first=o_first.cpu().numpy(),
                   free_delta=free_before - free_after)
    return rc, err.value.decode(), out


def _reference_tick(rsb, x, y, z, ch, enabled, src, power, pkt_channel):
    with rsb.Medium("udgm") as m:
        m.set_nodes(x, y, z, channel=ch, enabled=enabled)
        return m.tick(src, power_dbm=power, channel=pkt_channel), m.cfg.transmission_range


def _same(got, ref):
    assert got["n_pairs"] == ref.n_pairs
    assert np.array_equal(got["tx"], ref.tx) and np.array_equal(got["rx"], ref.rx)
    assert np.array_equal(got["outcome"], ref.outcome)
    assert np.array_equal(got["rssi"], ref.rssi_dbm)                   # float RSSI = (float) tx power on both sides: bit-exact
    first = np.searchsorted(ref.tx, np.arange(ref.n_tx + 1)).astype(np.int32)
    assert np.array_equal(got["tx_first"], first)


def test_static_tick_equals_rmb_tick_on_two_calls_in_a_row():
    import torch
    import radio_sim_b200 as rsb
    from radio_sim_b200 import _abi as A
    lib = A.load()
    # call 1: 5 000 nodes x 300 packets, two channels, some radios off, mixed power, per-packet channel override,
    # a node that sends twice
    n1 = 5000
    x1, y1, z1, ch1 = random_deployment(301, n1, 700.0, channels=2, z_spread=20.0)
    en1 = np.ones(n1, dtype=np.uint8); en1[::17] = 0
    src1 = pick_sources(301, n1, 300); src1[7] = src1[3]
    pw1 = np.where(np.arange(300) % 3 == 0, -7.0, 0.0)
    pc1 = ch1[src1].copy(); pc1[::11] = 11
    ref1, R = _reference_tick(rsb, x1, y1, z1, ch1, en1, src1, pw1, pc1)
    # call 2: smaller and different (1 200 nodes x 40 packets, planar, channel taken from the source)
    n2 = 1200
    x2, y2, _, ch2 = random_deployment(302, n2, 300.0, channels=3)
    src2 = pick_sources(302, n2, 40)
    pw2 = np.full(40, -3.3)
    ref2, _ = _reference_tick(rsb, x2, y2, None, ch2, None, src2, pw2, None)
    assert ref1.n_pairs > 1000 and ref2.n_pairs > 50

    rc, err, got1 = _static_tick(lib, torch, x1, y1, z1, ch1, en1, src1, pw1, pc1, R, ref1.n_pairs + 10)
    assert rc == 0, err
    rc, err, got2 = _static_tick(lib, torch, x2, y2, None, ch2, None, src2, pw2, None, R, ref2.n_pairs)
    assert rc == 0, err
    _same(got1, ref1)
    _same(got2, ref2)
    assert got2["free_delta"] <= 0                       # nothing was allocated on the device by the second call
    # and the first tick once more over the second's leftovers
    rc, err, got3 = _static_tick(lib, torch, x1, y1, z1, ch1, en1, src1, pw1, pc1, R, ref1.n_pairs)
    assert rc == 0, err
    _same(got3, ref1)
    assert got3["free_delta"] <= 0


def test_static_tick_limits_and_short_capacity():
    import torch
    import radio_sim_b200 as rsb
    from radio_sim_b200 import _abi as A
    lib = A.load()
    n = 3000
    x, y, z, ch = random_deployment(303, n, 400.0)
    src = pick_sources(303, n, 60)
    pw = np.zeros(60)
    ref, R = _reference_tick(rsb, x, y, z, ch, None, src, pw, None)
    # fewer rows than the tick has: the count is the tick's, the rows that fit are the first rows
    cap = ref.n_pairs // 2
    rc, err, got = _static_tick(lib, torch, x, y, z, ch, None, src, pw, None, R, cap, want_tx=False)
    assert rc == 0, err
    assert got["n_pairs"] == ref.n_pairs and len(got["rx"]) == cap
    assert np.array_equal(got["rx"], ref.rx[:cap]) and np.array_equal(got["outcome"], ref.outcome[:cap])
    # an empty tick (what a caller issues before a graph capture to get the per-device set-up done)
    rc, err, got = _static_tick(lib, torch, x, y, z, ch, None, np.zeros(0, dtype=np.int32), np.zeros(0), None, R, 0)
    assert rc == 0 and got["n_pairs"] == 0 and list(got["tx_first"]) == [0]
    # beyond the fixed scratch: n_tx * ceil(n / 256) > RMB_STATIC_MAX_CELLS is an error, not a wrong result
    big_t = (1 << 19) // ((n + 255) // 256) + 1
    rc, err, got = _static_tick(lib, torch, x, y, z, ch, None, np.zeros(big_t, dtype=np.int32), np.zeros(big_t), None, R, 16)
    assert rc == A.ERR_CAPACITY and "RMB_STATIC_MAX_CELLS" in err and got is None
    rc, err, got = _static_tick(lib, torch, x, y, z, ch, None, src, None, None, R, 16)       # no power column
    assert rc == A.ERR_INVALID and "NULL" in err
